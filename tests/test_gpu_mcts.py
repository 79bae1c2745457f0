"""GPU tests of the drop-in Evaluator behind MCTS (modelled on the reference's own tests:
python/algorithms/mcts_test.py:70-153, alpha_zero/evaluator_test.py:35-104)."""
import numpy as np
import pytest

from oracle_lib import Oracle

pytestmark = pytest.mark.gpu
UCT_C = 2.0


@pytest.fixture(scope="module")
def bsg():
  import open_spiel_1_b200 as m
  return m


def test_evaluate_matches_oracle_rollouts(bsg):
  """evaluate() = mean of the oracle's replays of the very playout ids it consumed."""
  oracle = Oracle()
  for game in ("tic_tac_toe", "connect_four", "breakthrough", "hex", "pentago", "othello", "y", "havannah", "go(board_size=9)"):
    ev = bsg.GpuRolloutEvaluator(game, n_rollouts=64, seed=123)
    g = bsg.load_game(game)
    state = g.new_initial_state()
    rng = np.random.RandomState(1)
    for _ in range(5):
      la = state.legal_actions()
      state.apply_action(la[rng.randint(len(la))])
    before = state.history()
    for call in range(3):
      v = ev.evaluate(state)
      assert state.history() == before               # evaluate must not mutate the state
      assert len(v) == 2 and v[0] + v[1] == 0         # zero-sum (evaluator_test.py)
      lo, hi = ev.last_playout_range
      _, o_ret, _, _ = oracle.replay_batch(game, [before], rollouts=hi - lo, seed=123,
                                           first_playout=lo)
      assert v[0] == o_ret.astype(np.float64).mean()
    pri = ev.prior(state)
    assert isinstance(pri, list)
    assert [a for a, _ in pri] == state.legal_actions()
    assert abs(sum(p for _, p in pri) - 1.0) < 1e-12
    ev.engine.close()


def test_evaluate_batch_equals_single_calls(bsg):
  g = bsg.load_game("connect_four")
  states = []
  rng = np.random.RandomState(3)
  for i in range(17):
    s = g.new_initial_state()
    for _ in range(i % 9):
      la = s.legal_actions()
      s.apply_action(la[rng.randint(len(la))])
    states.append(s)
  ev_a = bsg.GpuRolloutEvaluator("connect_four", n_rollouts=32, seed=9)
  ev_b = bsg.GpuRolloutEvaluator(ev_a.engine, n_rollouts=32, seed=9)
  batch = ev_a.evaluate_batch(states)
  single = np.stack([ev_b.evaluate(s) for s in states])
  assert np.array_equal(batch, single)


def test_mcts_plays_tic_tac_toe_games(bsg):
  """mcts_test.py:70-81: two bots sharing one evaluator play a full game; returns sum to 0."""
  game = bsg.load_game("tic_tac_toe")
  rs = np.random.RandomState(42)
  ev = bsg.GpuRolloutEvaluator("tic_tac_toe", n_rollouts=20, seed=42)
  bots = [bsg.MCTSBot(game, UCT_C, 100, ev, random_state=rs),
          bsg.MCTSBot(game, UCT_C, 100, ev, random_state=rs)]
  state = game.new_initial_state()
  while not state.is_terminal():
    state.apply_action(bots[state.current_player()].step(state))
  assert sum(state.returns()) == 0


def _search(bsg, history, leaf_batch=1):
  game = bsg.load_game("tic_tac_toe")
  state = game.new_initial_state()
  for a in history:
    state.apply_action(a)
  rs = np.random.RandomState(42)
  ev = bsg.GpuRolloutEvaluator("tic_tac_toe", n_rollouts=20, seed=42)
  bot = bsg.MCTSBot(game, UCT_C, 10000, ev, random_state=rs, leaf_batch=leaf_batch)
  return bot.mcts_search(state), state


@pytest.mark.parametrize("leaf_batch", [1, 8])
def test_solver_known_answers(bsg, leaf_batch):
  """mcts_test.py:115-153: draw / loss / win positions are proven by the solver."""
  root, state = _search(bsg, [4, 0, 8], leaf_batch)          # x(1,1) o(0,0) x(2,2)
  assert root.outcome[root.player] == 0
  assert all(c.outcome[c.player] <= 0 for c in root.children)
  best = root.best_child()
  assert best.outcome[best.player] == 0 and best.action in (2, 6)
  root, state = _search(bsg, [4, 0, 8, 1, 2], leaf_batch)    # all moves lose
  assert root.outcome[root.player] == -1
  assert all(c.outcome[c.player] == -1 for c in root.children)
  root, state = _search(bsg, [1, 8], leaf_batch)             # x(0,1) o(2,2): x(0,2) wins
  assert root.outcome[root.player] == 1
  best = root.best_child()
  assert best.outcome[best.player] == 1 and best.action == 2


def test_best_child_ordering(bsg):
  """mcts_test.py:155-216: proven win > unproven visits > proven loss; ties by count/reward."""
  def node(count, reward, outcome):
    n = bsg.SearchNode(0, 0, 1)
    n.explore_count, n.total_reward = count, reward
    n.outcome = None if outcome is None else [outcome, -outcome]
    return n
  def best(children):
    root = bsg.SearchNode(None, 0, 1)
    root.children = children
    return root.best_child()
  a, b = node(10, 5, 1), node(100, 90, None)
  assert best([a, b]) is a                  # proven win beats a promising unproven move
  a, b = node(10, 1, 0), node(20, 5, None)
  assert best([a, b]) is b                  # more visits beat a proven draw with fewer
  a, b = node(30, 1, 0), node(20, 5, None)
  assert best([a, b]) is a
  a, b = node(50, -40, -1), node(3, 0, None)
  assert best([a, b]) is b                  # anything unproven beats a proven loss
  a, b = node(10, 4, None), node(10, 5, None)
  assert best([a, b]) is b                  # equal visits: higher reward


def _root_stats(root):
  return sorted((c.action, c.explore_count, c.total_reward,
                 None if c.outcome is None else tuple(c.outcome)) for c in root.children)


@pytest.mark.parametrize("game,moves,leaf_batch", [
    ("tic_tac_toe", 0, 1), ("tic_tac_toe", 2, 8), ("connect_four", 6, 1), ("connect_four", 9, 16),
    ("breakthrough", 10, 8), ("hex", 30, 4), ("pentago", 8, 8), ("othello", 12, 8), ("y", 30, 4), ("havannah", 30, 4), ("go(board_size=9)", 30, 4)])
def test_native_tree_equals_python_mirror(bsg, game, moves, leaf_batch):
  """The native host tree (csrc/bsg_mcts.cuh) builds the same tree as the Python mirror — the
  specification, itself a restatement of mcts.py:199-443 — visit for visit: same shuffles
  (PhiloxShuffle), same playout ids, same UCT arithmetic, same solver backups; also when a
  search is resumed."""
  g = bsg.load_game(game)
  state = g.new_initial_state()
  rng = np.random.RandomState(5)
  for _ in range(moves):
    la = state.legal_actions()
    state.apply_action(la[rng.randint(len(la))])
  sims, n_roll = 300, 8
  eng = bsg.Engine(game, 0)
  ev = bsg.GpuRolloutEvaluator(eng, n_rollouts=n_roll, seed=77)
  py = bsg.MCTSBot(g, UCT_C, sims, ev, solve=True, random_state=bsg.PhiloxShuffle(13),
                   leaf_batch=leaf_batch)
  nat = bsg.NativeMCTSBot(g, UCT_C, sims, eng, n_rollouts=n_roll, solve=True, seed=77,
                          shuffle_seed=13, leaf_batch=leaf_batch)
  r_py = py.mcts_search(state)
  r_nat = nat.mcts_search(state)
  assert r_nat.explore_count == r_py.explore_count
  assert _root_stats(r_nat) == _root_stats(r_py)
  assert nat.next_playout == ev.next_playout
  assert [c.action for c in r_nat.children] == [c.action for c in r_py.children]   # same shuffle
  assert r_nat.best_child().action == r_py.best_child().action
  # resume both with 100 more simulations
  r_py = py.mcts_search(state, root=r_py, simulations=100)
  r_nat = nat.mcts_search(state, root=r_nat, simulations=100)
  assert _root_stats(r_nat) == _root_stats(r_py)
  nat.close()
  eng.close()


def test_native_tree_solver_known_answers(bsg):
  """mcts_test.py:115-153 on the native tree: proven win, proven loss, proven draw."""
  g = bsg.load_game("tic_tac_toe")
  eng = bsg.Engine("tic_tac_toe", 0)
  # x to move after x(4) o(1) x(0) o(8): x wins (e.g. 6 or 3 threatens twice); solver proves +1
  state = g.new_initial_state()
  for a in [4, 1, 0, 8]:
    state.apply_action(a)
  nat = bsg.NativeMCTSBot(g, UCT_C, 5000, eng, n_rollouts=1, solve=True, seed=1, shuffle_seed=2,
                          leaf_batch=4)
  root = nat.mcts_search(state)
  assert root.outcome is not None and root.outcome[state.current_player()] == 1.0
  best = root.best_child()
  assert best.outcome is not None and best.outcome[best.player] == 1.0
  # the python mirror agrees on the proven value
  ev = bsg.GpuRolloutEvaluator(eng, n_rollouts=1, seed=1)
  py = bsg.MCTSBot(g, UCT_C, 5000, ev, solve=True, random_state=bsg.PhiloxShuffle(2), leaf_batch=4)
  r_py = py.mcts_search(state)
  assert r_py.outcome == root.outcome
  nat.close()
  eng.close()


class _TinyModel(object):
  """Deterministic stand-in for alpha_zero/model.py: value = tanh(w_v . obs), policy = softmax
  of W_p obs over the legal actions; float64 so host and device agree to rounding."""

  def __init__(self, obs_size, num_actions, device):
    import torch
    g = torch.Generator().manual_seed(4)
    self.wv = (torch.randint(-3, 4, (obs_size,), generator=g).double() / 16).to(device)
    self.wp = (torch.randint(-3, 4, (num_actions, obs_size), generator=g).double() / 16).to(device)
    self.calls = 0

  def inference(self, obs, mask):
    import torch
    self.calls += 1
    x = obs.reshape(obs.shape[0], -1).double().to(self.wv.device)
    mask = torch.as_tensor(mask).to(self.wv.device).bool()
    value = torch.tanh(x @ self.wv).reshape(-1, 1)
    logits = (x @ self.wp.T).masked_fill(~mask, -1e30)
    return value, torch.softmax(logits, dim=1)


@pytest.mark.parametrize("game", ["tic_tac_toe", "connect_four", "breakthrough", "hex", "pentago", "othello", "y", "havannah", "go(board_size=9)"])
def test_alpha_zero_evaluator_matches_host_observations(bsg, game):
  """AlphaZeroEvaluator (evaluator.py:38-83 of the reference) over the GPU observation writer:
  same value / prior as feeding the model the host mirror's observation_tensor() and
  legal_actions_mask(), zero-sum evaluate, priors over the legal actions only, cache hits."""
  import torch
  g = bsg.load_game(game)
  dev = torch.device("cuda", 0)
  model = _TinyModel(g.observation_tensor_size(), g.num_distinct_actions(), dev)
  host_model = _TinyModel(g.observation_tensor_size(), g.num_distinct_actions(), torch.device("cpu"))
  ev = bsg.AlphaZeroEvaluator(g, model)
  rng = np.random.RandomState(2)
  states = []
  for i in range(9):
    s = g.new_initial_state()
    for _ in range(2 * i):
      if s.is_terminal():
        break
      la = s.legal_actions()
      s.apply_action(la[rng.randint(len(la))])
    if not s.is_terminal():
      states.append(s)
  batch = ev.evaluate_batch(states)
  for i, s in enumerate(states):
    obs = torch.tensor(np.asarray(s.observation_tensor(), np.float64)).reshape(1, -1)
    mask = torch.tensor(np.asarray(s.legal_actions_mask(), np.bool_)).reshape(1, -1)
    v, p = host_model.inference(obs, mask)
    e = ev.evaluate(s)
    assert abs(e[0] - float(v[0, 0])) < 1e-9 and e[0] + e[1] == 0
    assert np.array_equal(batch[i], e)
    pri = ev.prior(s)
    assert [a for a, _ in pri] == s.legal_actions()
    assert abs(sum(x for _, x in pri) - 1.0) < 1e-9
    for a, x in pri:
      assert abs(x - float(p[0, a])) < 1e-9
  info = ev.cache_info()
  assert model.calls == 1 and info["hits"] > 0     # one batched model call, the rest cached
  # drives the Python MCTS like the reference's AlphaZero actor (alpha_zero.py:196-228)
  bot = bsg.MCTSBot(g, 1.5, 30, ev, solve=False, random_state=np.random.RandomState(0),
                    child_selection_fn=bsg.SearchNode.puct_value)
  action = bot.step(g.new_initial_state())
  assert action in g.new_initial_state().legal_actions()
  ev.engine.close()


def _root_table(root):
  return sorted((c.action, c.explore_count, round(c.total_reward * 64)) for c in root.children)


def test_trees_per_rank_is_deterministic_and_equals_virtual_ranks(bsg):
  """K trees per rank (one host thread + one engine each): the merged root statistics are the
  same whether the trees run concurrently or one after the other, and — for a single merge
  step — equal the sum of the K single-tree searches that virtual ranks 0..K-1 would run
  alone (same shuffle seeds, playout-id ranges and simulation shares)."""
  from open_spiel_1_b200.parallel import RootParallelMCTS, shard_range
  game = bsg.load_game("connect_four")
  state = game.new_initial_state()
  for a in (3, 3, 2):
    state.apply_action(a)
  K, sims, n_roll = 3, 3000, 64

  def run(threads, sims_per_step):
    ev = bsg.GpuRolloutEvaluator("connect_four", n_rollouts=n_roll, seed=21)
    rp = RootParallelMCTS(game, UCT_C, sims, ev, sims_per_step=sims_per_step, solve=True, seed=9,
                          leaf_batch=32, native=True, trees_per_rank=K, threads=threads)
    root = rp.mcts_search(state)
    return _root_table(root), rp

  multi_step, rp = run(True, 256)
  assert rp.n_steps == 4 and len(rp.bots) == K
  assert multi_step == run(False, 256)[0] == run(True, 256)[0]
  assert sum(v for _, v, _ in multi_step) == sims - K   # each tree's first visit evaluates the root itself

  one_step, _ = run(True, None)
  total = {}
  for v in range(K):
    lo, hi = shard_range(sims, v, K)
    eng = bsg.Engine("connect_four", 0)
    bot = bsg.NativeMCTSBot(game, UCT_C, hi - lo, eng, n_rollouts=n_roll, solve=True, seed=21,
                            shuffle_seed=9 + v, leaf_batch=32)
    bot.next_playout = v << 40
    for a, n, r in _root_table(bot.mcts_search(state)):
      t = total.setdefault(a, [0, 0])
      t[0] += n
      t[1] += r
    bot.close()
    eng.close()
  assert one_step == sorted((a, n, r) for a, (n, r) in total.items())


def test_tournament_from_gpu_roots(bsg):
  """evaluate_bots / play_match (algorithms/evaluate_bots.cc:24-62) with GPU-evaluated search
  bots, from roots the playout kernel generated: a native-tree MCTS bot beats uniform-random
  play at connect_four from both seats."""
  from open_spiel_1_b200.roots import make_roots
  game = bsg.load_game("connect_four")
  eng = bsg.Engine("connect_four", 0)
  _, hists = make_roots(eng, 24, 0, 4, seed=5, want_histories=True)
  score = 0.0
  for seat in (0, 1):
    mcts_bot = bsg.NativeMCTSBot(game, UCT_C, 300, eng, n_rollouts=16, seed=3, shuffle_seed=seat,
                                 leaf_batch=8)
    rnd = bsg.UniformRandomBot(1 - seat, np.random.RandomState(seat))
    bots = [mcts_bot, rnd] if seat == 0 else [rnd, mcts_bot]
    roots = []
    for h in hists:
      s = game.new_initial_state()
      for a in h:
        s.apply_action(a)
      roots.append(s)
    mean, results = bsg.play_match(roots, bots)
    assert all(r.is_terminal() for r in roots) and (results.sum(axis=1) == 0).all()
    score += mean[seat]
    mcts_bot.close()
  assert score > 1.2          # wins the large majority from either seat
  eng.close()


@pytest.mark.parametrize("game", ["hex", "y"])
def test_value_only_evaluator_returns_the_same_values(bsg, game):
  """GpuRolloutEvaluator(value_only=True): evaluate / evaluate_batch give exactly the values of
  the exact evaluator for the same playout ids (chunked host path included); games without the
  property refuse."""
  g = bsg.load_game(game)
  rng = np.random.RandomState(4)
  states = []
  for i in range(40):
    s = g.new_initial_state()
    for _ in range(3 * i):
      if s.is_terminal():
        break
      la = s.legal_actions()
      s.apply_action(la[rng.randint(len(la))])
    if not s.is_terminal():
      states.append(s)
  ev_a = bsg.GpuRolloutEvaluator(game, n_rollouts=64, seed=5)
  ev_b = bsg.GpuRolloutEvaluator(game, n_rollouts=64, seed=5, value_only=True)
  for s in states[:6]:
    assert np.array_equal(ev_a.evaluate(s), ev_b.evaluate(s))
  assert np.array_equal(ev_a.evaluate_batch(states), ev_b.evaluate_batch(states))
  big = states * 3000                       # > 2 MiB of roots: several chunks on two streams
  ev_a.n_rollouts = ev_b.n_rollouts = 1
  assert np.array_equal(ev_a.evaluate_batch(big), ev_b.evaluate_batch(big))
  assert ev_b.total_steps >= ev_a.total_steps
  ev_a.engine.close(); ev_b.engine.close()
  with pytest.raises(RuntimeError):
    bsg.GpuRolloutEvaluator("breakthrough", n_rollouts=4, value_only=True)


def test_native_tree_over_a_value_only_engine_builds_the_same_tree(bsg):
  """bsg_mcts_* evaluates its leaves through bsg_playouts_host, which honours
  bsg_engine_set_value_only: same tree, visit for visit."""
  g = bsg.load_game("hex")
  state = g.new_initial_state()
  for a in (60, 48, 61):
    state.apply_action(a)
  roots = []
  for vo in (False, True):
    eng = bsg.Engine("hex", 0)
    if vo:
      eng.set_value_only(True)
    bot = bsg.NativeMCTSBot(g, UCT_C, 400, eng, n_rollouts=8, solve=True, seed=77, shuffle_seed=13, leaf_batch=8)
    roots.append(_root_stats(bot.mcts_search(state)))
    bot.close()
    eng.close()
  assert roots[0] == roots[1]


def test_batched_search_from_a_deserialized_root(bsg):
  """A connect_four root read from a board string has no history (as in the reference,
  connect_four.cc:229-279); the leaves of a batched search below it must still be evaluated at
  THEIR positions: the search equals the one from the same position reached by its history."""
  game = bsg.load_game("connect_four")
  played = game.new_initial_state()
  for a in (3, 3, 2, 4, 1, 5, 3, 2):
    played.apply_action(a)
  loaded = game.deserialize_state(played.serialize())
  assert loaded.history() == [] and np.array_equal(loaded.packed(), played.packed())
  roots = []
  for root_state, lb in ((played, 1), (played, 8), (loaded, 8), (loaded, 1)):
    ev = bsg.GpuRolloutEvaluator("connect_four", n_rollouts=16, seed=5)
    bot = bsg.MCTSBot(game, UCT_C, 200, ev, solve=True, random_state=bsg.PhiloxShuffle(11),
                      leaf_batch=lb)
    root = bot.mcts_search(root_state)
    roots.append((lb, sorted((c.action, c.explore_count, c.total_reward) for c in root.children)))
    ev.engine.close()
  assert roots[0][1] == roots[3][1]      # leaf_batch 1: history vs board string
  assert roots[1][1] == roots[2][1]      # leaf_batch 8: history vs board string
