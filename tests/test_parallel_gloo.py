"""CPU tests of the N > 1 path (world_size 2, gloo): shard ranges, the totals reduction and the
root-parallel MCTS merge (parallel.py).  The playout kernels need a GPU, so the evaluator here
is an oracle-backed CHECKER stand-in with the same playout-id contract as
GpuRolloutEvaluator (test infrastructure only)."""
import os
import socket
import sys

import numpy as np
import pytest

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, REPO)
sys.path.insert(0, os.path.join(REPO, "tests"))


def _free_port():
  s = socket.socket()
  s.bind(("127.0.0.1", 0))
  port = s.getsockname()[1]
  s.close()
  return port


class OracleRolloutEvaluator:
  """Mean of the oracle's replays of playout ids [next, next + n) from state.history()."""

  def __init__(self, game, n_rollouts, seed):
    from oracle_lib import Oracle
    self.oracle, self.game = Oracle(), game
    self.n_rollouts, self.seed, self.next_playout = n_rollouts, seed, 0

  def evaluate(self, state):
    _, ret, _, _ = self.oracle.replay_batch(self.game, [state.history()], rollouts=self.n_rollouts,
                                            seed=self.seed, first_playout=self.next_playout)
    self.next_playout += self.n_rollouts
    v = float(ret.astype(np.int64).sum()) / self.n_rollouts
    return np.array([v, -v])

  def evaluate_batch(self, states):
    return np.stack([self.evaluate(s) for s in states])

  def prior(self, state):
    la = state.legal_actions(state.current_player())
    return [(a, 1.0 / len(la)) for a in la]


def _search(rank, world, game_name, history, sims, sims_per_step, rollouts, leaf_batch=1):
  import open_spiel_1_b200 as bsg
  from open_spiel_1_b200.parallel import RootParallelMCTS
  game = bsg.load_game(game_name)
  state = game.new_initial_state()
  for a in history:
    state.apply_action(a)
  ev = OracleRolloutEvaluator(game_name, rollouts, seed=77)
  rp = RootParallelMCTS(game, 2.0, sims, ev, sims_per_step=sims_per_step, seed=5,
                        leaf_batch=leaf_batch)
  root = rp.mcts_search(state)
  stats = sorted((c.action, c.explore_count, int(round(c.total_reward * rollouts)),
                  None if c.outcome is None else tuple(c.outcome)) for c in root.children)
  # the move step_with_policy would pick: best sort_key, ties to the lowest action id
  action = max(sorted(root.children, key=lambda c: c.action), key=lambda c: c.sort_key()).action
  return {"stats": stats, "action": action, "all_reduces": rp.all_reduces,
          "local_sims": rp.local_simulations, "n_steps": rp.n_steps}


def _worker(rank, world, port, q):
  import torch.distributed as dist
  from open_spiel_1_b200.parallel import shard_range, all_reduce_int64
  dist.init_process_group("gloo", init_method="tcp://127.0.0.1:%d" % port, rank=rank,
                          world_size=world)
  out = {}
  # (1) totals reduction over disjoint shards: int64 sums, order-independent
  lo, hi = shard_range(1001, rank, world)
  mine = np.array([hi - lo, sum(range(lo, hi)), -(hi - lo), 0], np.int64)
  out["totals"] = all_reduce_int64(mine).tolist()
  # (2) root-parallel search of connect_four, 3 merges
  out["c4"] = _search(rank, world, "connect_four", [3, 3, 2], 240, 40, 4)
  # (3) a solver case: ttt position where x wins at once (mcts_test.py:115-127 shape)
  out["ttt"] = _search(rank, world, "tic_tac_toe", [4, 0, 2, 1], 200, 50, 1)
  # (4) batched leaves inside each rank
  out["c4_batched"] = _search(rank, world, "connect_four", [3], 128, 32, 2, leaf_batch=4)
  q.put((rank, out))
  dist.barrier()
  dist.destroy_process_group()


@pytest.fixture(scope="module")
def two_ranks():
  import torch.multiprocessing as mp
  ctx = mp.get_context("spawn")
  q = ctx.Queue()
  port = _free_port()
  procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
  for p in procs:
    p.start()
  res = {}
  for _ in procs:
    rank, out = q.get(timeout=240)
    res[rank] = out
  for p in procs:
    p.join(timeout=60)
    assert p.exitcode == 0
  return res


def test_shard_range_partitions():
  from open_spiel_1_b200.parallel import shard_range
  for n in (0, 1, 7, 16, 1001, 1 << 24):
    for world in (1, 2, 3, 8):
      cuts = [shard_range(n, r, world) for r in range(world)]
      assert cuts[0][0] == 0 and cuts[-1][1] == n
      for a, b in zip(cuts, cuts[1:]):
        assert a[1] == b[0]
      sizes = [hi - lo for lo, hi in cuts]
      assert max(sizes) - min(sizes) <= 1
  with pytest.raises(ValueError):
    shard_range(10, 2, 2)


def test_totals_all_reduce(two_ranks):
  want = [1001, sum(range(1001)), -1001, 0]
  assert two_ranks[0]["totals"] == want and two_ranks[1]["totals"] == want


def test_root_parallel_merge_is_identical_on_all_ranks(two_ranks):
  for key, sims, steps in (("c4", 240, 3), ("c4_batched", 128, 2)):
    a, b = two_ranks[0][key], two_ranks[1][key]
    assert a["stats"] == b["stats"]            # same global visits / integer rewards everywhere
    assert a["action"] == b["action"]          # hence the same move on every rank
    assert a["all_reduces"] == b["all_reduces"] == steps == a["n_steps"]
    assert a["local_sims"] + b["local_sims"] == sims
    # every simulation of every rank is counted; each rank's first one evaluates the root
    # itself and visits no child (mcts.py:347-405)
    assert sum(s[1] for s in a["stats"]) == sims - 2


def test_root_parallel_solver_proofs_travel(two_ranks):
  a, b = two_ranks[0]["ttt"], two_ranks[1]["ttt"]
  assert a["stats"] == b["stats"] and a["action"] == b["action"]
  # x (player 0) to move with 4,2 on the anti-diagonal: playing 6 wins on the spot
  assert a["action"] == 6
  proven = {s[0]: s[3] for s in a["stats"] if s[3] is not None}
  assert proven.get(6) == (1.0, -1.0)


def test_world_size_one_is_a_noop():
  res = _search(0, 1, "connect_four", [3, 3, 2], 60, 20, 2)
  assert res["all_reduces"] == 3 and sum(s[1] for s in res["stats"]) == 60 - 1
