"""The reference's python/algorithms/mcts.py, UNMODIFIED, against this repo's host mirror (CPU).

oracle/Makefile builds a reduced `pyspiel` module over the unmodified reference C++
(oracle/pyspiel_min.cc) and byte-compiles the reference's mcts.py from where it lies.  With
those, the reference's own MCTSBot is the specification the mirror `mcts.MCTSBot` (and through
it the native tree, tests/test_gpu_mcts.py) is held to: same evaluator, same random streams ->
the same tree, visit for visit.  Skipped when oracle/_ref/ has not been built (the reference
tree is needed once to build it)."""
import numpy as np
import pytest

from oracle_lib import reference_python

UCT_C = 2.0
GAMES = ["tic_tac_toe", "connect_four", "breakthrough", "hex(board_size=11)", "pentago", "othello", "y", "havannah", "go(board_size=9)"]


@pytest.fixture(scope="module")
def refpy():
  mods = reference_python()
  if mods is None:
    pytest.skip("oracle/_ref/pyspiel*.so or ref_mcts.bytecode not built")
  return mods


def tree_dump(node, depth=0, max_depth=3, out=None):
  """(depth, action, player, visits, reward, outcome) of every node down to max_depth, children
  in list order (the order the bot's random_state shuffled them into)."""
  out = [] if out is None else out
  out.append((depth, node.action, node.player, node.explore_count, round(node.total_reward, 9),
              None if node.outcome is None else tuple(node.outcome)))
  if depth < max_depth:
    for ch in node.children:
      tree_dump(ch, depth + 1, max_depth, out)
  return out


def advance(state, moves, rng):
  for _ in range(moves):
    if state.is_terminal():
      break
    la = state.legal_actions()
    state.apply_action(la[rng.randint(len(la))])
  return state


@pytest.mark.parametrize("game", GAMES)
def test_mirror_bot_builds_the_reference_bots_tree(refpy, game):
  """Reference MCTSBot on pyspiel states vs the mirror MCTSBot on mirror states, both driven
  by the reference's own RandomRolloutEvaluator with equal seeds."""
  pyspiel, ref_mcts = refpy
  import open_spiel_1_b200 as bsg
  rng = np.random.RandomState(2)
  rg, mg = pyspiel.load_game(game), bsg.load_game(game)
  for trial in range(3):
    rs, ms = rg.new_initial_state(), mg.new_initial_state()
    depth = [0, 3, 6] if game == "tic_tac_toe" else [0, 6, 15]
    advance(rs, depth[trial], rng)
    while rs.is_terminal():                      # mcts_search needs a decision node
      rs = advance(rg.new_initial_state(), depth[trial], rng)
    for a in rs.history():
      ms.apply_action(a)
    assert str(ms) == str(rs)
    sims = 300 if game.startswith("hex") else (100 if game in ("y", "havannah") or game.startswith("go") else 600)   # y: ~170 steps per rollout
    r_bot = ref_mcts.MCTSBot(rg, UCT_C, sims, ref_mcts.RandomRolloutEvaluator(2, np.random.RandomState(7)),
                             solve=True, random_state=np.random.RandomState(3))
    m_bot = bsg.MCTSBot(mg, UCT_C, sims, ref_mcts.RandomRolloutEvaluator(2, np.random.RandomState(7)),
                        solve=True, random_state=np.random.RandomState(3))
    r_root, m_root = r_bot.mcts_search(rs), m_bot.mcts_search(ms)
    assert tree_dump(m_root) == tree_dump(r_root)
    assert m_root.best_child().action == r_root.best_child().action
    assert m_bot.step(ms) == r_bot.step(rs)        # fresh searches, streams still aligned


def test_reference_known_answers_hold_for_both_bots(refpy):
  """python/algorithms/mcts_test.py:115-153 (solve draw / loss / win) on the reference bot and
  on the mirror."""
  pyspiel, ref_mcts = refpy
  import open_spiel_1_b200 as bsg
  cases = [([4, 0, 8], 0, (2, 6)), ([4, 0, 8, 1, 2], -1, None), ([1, 8], 1, (2,))]
  for history, value, best_in in cases:
    for mod, load in ((ref_mcts, pyspiel.load_game), (bsg, bsg.load_game)):
      game = load("tic_tac_toe")
      state = game.new_initial_state()
      for a in history:
        state.apply_action(a)
      ev = ref_mcts.RandomRolloutEvaluator(20, np.random.RandomState(0))
      bot = mod.MCTSBot(game, UCT_C, 10000, ev, random_state=np.random.RandomState(0))
      root = bot.mcts_search(state)
      assert root.outcome[root.player] == value
      if best_in:
        assert root.best_child().action in best_in


def test_mirror_states_answer_like_pyspiel_states(refpy):
  """The pyspiel.State surface mcts.py relies on, method by method, along random games."""
  pyspiel, _ = refpy
  import open_spiel_1_b200 as bsg
  rng = np.random.RandomState(5)
  for game in GAMES:
    rg, mg = pyspiel.load_game(game), bsg.load_game(game)
    assert str(mg) == str(rg) and mg.max_utility() == rg.max_utility()
    assert mg.num_distinct_actions() == rg.num_distinct_actions()
    assert mg.max_game_length() == rg.max_game_length()
    assert list(mg.observation_tensor_shape()) == list(rg.observation_tensor_shape())
    assert mg.get_type().reward_model.name == rg.get_type().reward_model.name
    assert mg.get_type().dynamics.name == rg.get_type().dynamics.name
    for _ in range(5):
      rs, ms = rg.new_initial_state(), mg.new_initial_state()
      while True:
        assert ms.is_terminal() == rs.is_terminal()
        assert ms.current_player() == rs.current_player()
        assert ms.returns() == rs.returns() and ms.history() == rs.history()
        assert ms.is_chance_node() == rs.is_chance_node() is False
        assert str(ms) == str(rs)
        if rs.is_terminal():
          assert ms.legal_actions() == rs.legal_actions() == []
          break
        cur = rs.current_player()
        assert ms.legal_actions() == rs.legal_actions()
        assert ms.legal_actions(cur) == rs.legal_actions(cur)
        assert ms.legal_actions(1 - cur) == rs.legal_actions(1 - cur) == []
        assert ms.legal_actions_mask() == rs.legal_actions_mask()
        for p in (0, 1):
          assert ms.observation_tensor(p) == rs.observation_tensor(p)
        la = rs.legal_actions()
        a = la[rng.randint(len(la))]
        assert str(ms.child(a)) == str(rs.child(a))
        rs.apply_action(a)
        ms.apply_action(a)
      with pytest.raises(RuntimeError):
        rs.apply_action(0)
      with pytest.raises(RuntimeError):
        ms.apply_action(0)


def test_evaluate_bots_mirror_plays_like_the_reference_loop(refpy):
  """evaluate_bots (python/algorithms/evaluate_bots.py:24-50): with equal bot seeds the mirror
  loop over mirror states and the same loop over pyspiel states play the same games (the
  reference's own file imports the full pyspiel, so its sequential branch is restated here)."""
  pyspiel, ref_mcts = refpy
  import open_spiel_1_b200 as bsg

  def reference_loop(state, bots):
    for bot in bots:
      bot.restart_at(state)
    while not state.is_terminal():
      current_player = state.current_player()
      action = bots[current_player].step(state)
      state.apply_action(action)
    return state.returns()

  for game in ("tic_tac_toe", "connect_four", "pentago"):
    rg, mg = pyspiel.load_game(game), bsg.load_game(game)
    total = np.zeros(2)
    for g in range(15):
      bots_m = [bsg.UniformRandomBot(0, np.random.RandomState(g)), bsg.UniformRandomBot(1, np.random.RandomState(100 + g))]
      bots_r = [bsg.UniformRandomBot(0, np.random.RandomState(g)), bsg.UniformRandomBot(1, np.random.RandomState(100 + g))]
      ms, rs = mg.new_initial_state(), rg.new_initial_state()
      got = bsg.evaluate_bots(ms, bots_m, np.random)
      want = reference_loop(rs, bots_r)
      assert got == want and ms.history() == rs.history() and ms.is_terminal()
      total += got
    assert total.sum() == 0
  # an MCTS bot (reference evaluator on the CPU) against a random bot through play_match
  mg = bsg.load_game("tic_tac_toe")
  ev = ref_mcts.RandomRolloutEvaluator(2, np.random.RandomState(1))
  bots = [bsg.MCTSBot(mg, UCT_C, 200, ev, random_state=np.random.RandomState(2)),
          bsg.UniformRandomBot(1, np.random.RandomState(3))]
  mean, results = bsg.play_match([mg.new_initial_state() for _ in range(10)], bots)
  assert results.shape == (10, 2) and mean[0] + mean[1] == 0 and mean[0] > 0   # x never loses to random play here
