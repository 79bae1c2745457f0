"""GPU-resident search forest (csrc/bsg_forest.cuh, forest.ForestMCTS) against the host-tree
root-parallel search it must equal: parallel.RootParallelMCTS(native=True, trees_per_rank=T),
whose trees are tied to the reference's mcts.py by tests/test_gpu_mcts.py and
tests/test_reference_python.py."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu
UCT_C = 2.0


@pytest.fixture(scope="module")
def bsg():
  import open_spiel_1_b200 as m
  return m


def table(root, scale):
  return sorted((c.action, c.explore_count, int(round(c.total_reward * scale)),
                 None if c.outcome is None else tuple(c.outcome)) for c in root.children)


@pytest.mark.parametrize("game,history,trees,sims,step,batch", [
    ("connect_four", [3, 3, 2], 6, 3000, 128, 16),
    ("connect_four", [], 4, 1001, None, 1),          # uneven shares, one merge, unbatched
    ("tic_tac_toe", [4, 0], 5, 4000, 200, 8),        # solver proofs travel through the merge
    ("pentago", [112, 3], 3, 600, 64, 8),
    ("hex", [60, 48], 3, 450, 50, 4),
    ("breakthrough", [], 3, 450, 75, 8),
    ("othello", [19, 18], 3, 450, 75, 8), ("y", [40, 41], 3, 240, 60, 4),
    ("havannah", [112, 97], 3, 300, 60, 4), ("go(board_size=9)", [40, 41, 81], 3, 240, 60, 4)])
def test_forest_equals_host_tree_root_parallel(bsg, game, history, trees, sims, step, batch):
  from open_spiel_1_b200.forest import ForestMCTS
  from open_spiel_1_b200.parallel import RootParallelMCTS
  g = bsg.load_game(game)
  state = g.new_initial_state()
  for a in history:
    state.apply_action(a)
  n_roll = 16
  eng = bsg.Engine(game, 0)
  ev = bsg.GpuRolloutEvaluator(eng, n_rollouts=n_roll, seed=21)
  host = RootParallelMCTS(g, UCT_C, sims, ev, sims_per_step=step, solve=True, seed=9,
                          leaf_batch=batch, native=True, trees_per_rank=trees)
  want = host.mcts_search(state)
  ev2 = bsg.GpuRolloutEvaluator(eng, n_rollouts=n_roll, seed=21)
  forest = ForestMCTS(g, UCT_C, sims, ev2, trees, sims_per_step=step, solve=True, seed=9,
                      leaf_batch=batch)
  got = forest.mcts_search(state)
  assert forest.n_steps == host.n_steps
  assert table(got, n_roll) == table(want, n_roll)
  assert got.explore_count == want.explore_count
  # every tree of the forest holds the merged statistics; their own deeper nodes differ
  assert table(forest.tree_root(trees - 1, state.current_player()), n_roll) == table(want, n_roll)
  pick = lambda root: max(sorted(root.children, key=lambda c: c.action), key=bsg.SearchNode.sort_key).action
  assert pick(got) == pick(want)
  assert forest.step(state) in state.legal_actions()     # a second search on the same forest
  assert forest.total_steps > 0
  forest.close()
  eng.close()


def test_single_forest_tree_equals_native_bot(bsg):
  from open_spiel_1_b200.forest import ForestMCTS
  g = bsg.load_game("connect_four")
  state = g.new_initial_state()
  eng = bsg.Engine("connect_four", 0)
  ev = bsg.GpuRolloutEvaluator(eng, n_rollouts=8, seed=5)
  forest = ForestMCTS(g, UCT_C, 700, ev, 1, solve=True, seed=13, leaf_batch=8)
  got = forest.mcts_search(state)
  bot = bsg.NativeMCTSBot(g, UCT_C, 700, eng, n_rollouts=8, solve=True, seed=5, shuffle_seed=13,
                          leaf_batch=8)
  want = bot.mcts_search(state)
  assert [(c.action, c.explore_count, c.total_reward) for c in got.children] == \
      [(c.action, c.explore_count, c.total_reward) for c in want.children]     # same shuffle order too
  assert forest.total_steps == bot.total_steps
  bot.close()
  forest.close()
  eng.close()


def test_large_forest_accounts_for_every_simulation(bsg):
  """2048 trees x 64 simulations: visits add up, every tree ends with the same merged root, and
  the whole search is a handful of kernel launches per round."""
  from open_spiel_1_b200.forest import ForestMCTS
  g = bsg.load_game("connect_four")
  state = g.new_initial_state()
  eng = bsg.Engine("connect_four", 0)
  ev = bsg.GpuRolloutEvaluator(eng, n_rollouts=32, seed=1)
  trees, per_tree = 2048, 64
  forest = ForestMCTS(g, UCT_C, trees * per_tree, ev, trees, sims_per_step=32, solve=True, seed=3,
                      leaf_batch=16)
  root = forest.mcts_search(state)
  assert sum(c.explore_count for c in root.children) == trees * per_tree - trees   # first visit = the root itself
  assert forest.all_reduces == 2
  info = forest.info()
  # per search step the rounds are enqueued back to back; the last backup kernel reports whether
  # any tree still has simulations left
  assert info["rounds"] <= 8 and info["launches"] == 3 * info["rounds"]
  t0 = table(root, 32)
  for v in (1, 777, trees - 1):
    assert table(forest.tree_root(v, 0), 32) == t0
  forest.close()
  eng.close()


@pytest.mark.parametrize("game,history", [("hex", [60, 48]), ("y", [40, 41])])
def test_forest_with_value_only_leaves_builds_the_same_trees(bsg, game, history):
  """GpuRolloutEvaluator(value_only=True) behind the forest: the leaves are evaluated by the
  value-only kernels (root_playout_base + value_only in one launch, skipped slots included) and
  every root statistic equals the exact forest's."""
  from open_spiel_1_b200.forest import ForestMCTS
  g = bsg.load_game(game)
  state = g.new_initial_state()
  for a in history:
    state.apply_action(a)
  roots = []
  for vo in (False, True):
    eng = bsg.Engine(game, 0)
    ev = bsg.GpuRolloutEvaluator(eng, n_rollouts=16, seed=21, value_only=vo)
    forest = ForestMCTS(g, UCT_C, 600, ev, 4, sims_per_step=50, solve=True, seed=9, leaf_batch=4)
    roots.append(table(forest.mcts_search(state), 16))
    info = forest.info()
    assert info["launches"] == (4 if vo else 3) * info["rounds"]
    forest.close()
    eng.close()
  assert roots[0] == roots[1]


def test_engine_side_merge_equals_the_torch_merge(bsg):
  """bsg_forest_merge (pack, all-reduce, install, base advance on the engine's stream with no
  host synchronisation) gives the root the caller-run merge gives, step for step, also when a
  forest is used for a second move; world size 1 has no communicator and no collective."""
  from open_spiel_1_b200.forest import ForestMCTS
  from open_spiel_1_b200 import _cabi as C
  g = bsg.load_game("connect_four")
  tabs = {}
  for merge in ("engine", "torch"):
    eng = bsg.Engine("connect_four", 0)
    assert C.lib.bsg_comm_ranks(eng._h) == 1
    ev = bsg.GpuRolloutEvaluator(eng, n_rollouts=16, seed=4)
    forest = ForestMCTS(g, UCT_C, 64 * 96, ev, 64, sims_per_step=24, solve=True, seed=2, leaf_batch=8,
                        merge=merge)
    state = g.new_initial_state()
    out = []
    for _ in range(3):                                  # three moves on one forest
      root = forest.mcts_search(state)
      out.append(table(root, 16))
      state.apply_action(forest.step_with_policy(state)[1])
    assert forest.all_reduces == 4 * 6                  # 96 / 24 steps, two searches per move
    if merge == "engine":
      assert forest.allreduce_stats()["count"] == 0     # no communicator, no collective
    tabs[merge] = out
    forest.close()
    eng.close()
  assert tabs["engine"] == tabs["torch"]
