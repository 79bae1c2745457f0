"""The drop-in proper: the reference's UNMODIFIED python/algorithms/mcts.py (byte-compiled into
oracle/_ref/ by oracle/Makefile) searching over the reference's own compiled games (the reduced
`pyspiel` module of oracle/pyspiel_min.cc), with this repo's GpuRolloutEvaluator passed as its
`evaluator` — exactly what INTEGRATION.md section 1 tells a maintainer to do.  Nothing here
reads /root/reference at run time; the prebuilt files travel to the GPU box."""
import numpy as np
import pytest

from oracle_lib import Oracle, reference_python

pytestmark = pytest.mark.gpu
UCT_C = 2.0
GAMES = ["tic_tac_toe", "connect_four", "breakthrough", "hex(board_size=11)", "pentago", "othello", "y", "havannah", "go(board_size=9)"]


@pytest.fixture(scope="module")
def refpy():
  mods = reference_python()
  if mods is None:
    pytest.skip("oracle/_ref/pyspiel*.so or ref_mcts.bytecode not built")
  return mods


@pytest.fixture(scope="module")
def bsg():
  import open_spiel_1_b200 as m
  return m


def advance(state, moves, rng):
  for _ in range(moves):
    la = state.legal_actions()
    state.apply_action(la[rng.randint(len(la))])
  return state


def root_table(root):
  return [(c.action, c.explore_count, round(c.total_reward, 9),
           None if c.outcome is None else tuple(c.outcome)) for c in root.children]


@pytest.mark.parametrize("game", GAMES)
def test_evaluator_accepts_pyspiel_states(refpy, bsg, game):
  """evaluate(pyspiel.State) = mean of the oracle's replays of the playout ids it consumed;
  prior(pyspiel.State) = uniform over state.legal_actions(); the state is not mutated."""
  pyspiel, _ = refpy
  oracle = Oracle()
  name = game.split("(")[0]
  g = pyspiel.load_game(game)
  ev = bsg.GpuRolloutEvaluator(str(g), n_rollouts=48, seed=5)    # Game::ToString is accepted
  state = advance(g.new_initial_state(), 4, np.random.RandomState(1))
  before, text = state.history(), str(state)
  for _ in range(2):
    v = ev.evaluate(state)
    lo, hi = ev.last_playout_range
    _, o_ret, _, _ = oracle.replay_batch(name, [before], rollouts=hi - lo, seed=5, first_playout=lo)
    assert v[0] == o_ret.astype(np.float64).mean() and v[0] + v[1] == 0
  assert state.history() == before and str(state) == text
  pri = ev.prior(state)
  assert [a for a, _ in pri] == state.legal_actions()
  assert all(p == 1.0 / len(pri) for _, p in pri)


@pytest.mark.parametrize("game", GAMES)
def test_reference_bot_with_gpu_evaluator_equals_mirror_bot(refpy, bsg, game):
  """Same evaluator seed and shuffle stream: the reference MCTSBot over pyspiel states and
  the mirror MCTSBot over mirror states visit the same nodes and read the same rollouts."""
  pyspiel, ref_mcts = refpy
  rg, mg = pyspiel.load_game(game), bsg.load_game(game)
  rs = advance(rg.new_initial_state(), 5, np.random.RandomState(4))
  ms = mg.new_initial_state()
  for a in rs.history():
    ms.apply_action(a)
  eng = bsg.Engine(game, 0)
  roots = []
  for mod, g, s in ((ref_mcts, rg, rs), (bsg, mg, ms)):
    ev = bsg.GpuRolloutEvaluator(eng, n_rollouts=32, seed=17)
    bot = mod.MCTSBot(g, UCT_C, 200, ev, solve=True, random_state=np.random.RandomState(6))
    roots.append(bot.mcts_search(s))
  assert root_table(roots[0]) == root_table(roots[1])
  assert roots[0].best_child().action == roots[1].best_child().action
  eng.close()


def test_reference_bot_solver_known_answers_with_gpu_evaluator(refpy, bsg):
  """python/algorithms/mcts_test.py:115-153 with the GPU evaluator behind the reference bot."""
  pyspiel, ref_mcts = refpy
  game = pyspiel.load_game("tic_tac_toe")
  ev = bsg.GpuRolloutEvaluator("tic_tac_toe", n_rollouts=20, seed=1)
  for history, value, best_in in [([4, 0, 8], 0, (2, 6)), ([4, 0, 8, 1, 2], -1, None), ([1, 8], 1, (2,))]:
    state = game.new_initial_state()
    for a in history:
      state.apply_action(a)
    bot = ref_mcts.MCTSBot(game, UCT_C, 10000, ev, random_state=np.random.RandomState(0))
    root = bot.mcts_search(state)
    assert root.outcome[root.player] == value
    if best_in:
      assert root.best_child().action in best_in


def test_reference_bots_play_a_connect_four_game(refpy, bsg):
  """Two reference bots sharing one GPU evaluator (as mcts_test.py:70-78 shares one CPU
  evaluator) play connect_four to the end; returns are zero-sum."""
  pyspiel, ref_mcts = refpy
  game = pyspiel.load_game("connect_four")
  ev = bsg.GpuRolloutEvaluator("connect_four", n_rollouts=64, seed=2)
  bots = [ref_mcts.MCTSBot(game, UCT_C, 60, ev, random_state=np.random.RandomState(p)) for p in (0, 1)]
  state = game.new_initial_state()
  while not state.is_terminal():
    state.apply_action(bots[state.current_player()].step(state))
  assert sum(state.returns()) == 0 and len(state.history()) >= 7
