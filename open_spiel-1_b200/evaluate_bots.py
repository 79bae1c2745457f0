"""Play bots against each other — host mirror of the reference's tournament loop.

`evaluate_bots(state, bots, rng)` is `open_spiel/python/algorithms/evaluate_bots.py:24-50`
(C++ twin `algorithms/evaluate_bots.cc:24-62`) for the games this engine supports (sequential,
deterministic, perfect information: there are no chance or simultaneous nodes to handle, and
meeting one raises).  `UniformRandomBot` mirrors `python/bots/uniform_random.py` /
`spiel_bots.cc:31-64`.  `play_match` is the driver for GPU-generated roots (SURVEY.md section
8f row 4): it runs one game per root and averages the returns, the bots being e.g. `MCTSBot`s
or `NativeMCTSBot`s whose leaf evaluations run on the GPU.
"""
import numpy as np


class UniformRandomBot(object):
  """Chooses uniformly among `state.legal_actions(player_id)`."""

  def __init__(self, player_id, rng):
    self._player_id = player_id
    self._rng = rng

  def player_id(self):
    return self._player_id

  def restart_at(self, state):
    pass

  def inform_action(self, state, player_id, action):
    pass

  def step_with_policy(self, state):
    legal_actions = state.legal_actions(self._player_id)
    if not legal_actions:
      return [], -1   # pyspiel.INVALID_ACTION
    p = 1.0 / len(legal_actions)
    policy = [(action, p) for action in legal_actions]
    action = self._rng.choice(legal_actions)
    return policy, int(action)

  def step(self, state):
    return self.step_with_policy(state)[1]


def evaluate_bots(state, bots, rng=None):
  """Plays `bots` (one per player) from `state` to the end; returns the terminal utility for
  each bot.  `state` is advanced in place, as in the reference."""
  del rng  # only chance nodes consume it in the reference; the supported games have none
  for bot in bots:
    bot.restart_at(state)
  while not state.is_terminal():
    if state.is_chance_node():
      raise RuntimeError("bsg: chance nodes are outside the supported games")
    current_player = state.current_player()
    action = bots[current_player].step(state)
    for i, bot in enumerate(bots):
      if i != current_player and hasattr(bot, "inform_action"):
        bot.inform_action(state, current_player, action)
    state.apply_action(action)
  return state.returns()


def play_match(roots, bots):
  """One game per root state (each is advanced in place); returns (mean returns [2], returns
  [n, 2])."""
  results = np.array([evaluate_bots(root, bots) for root in roots], np.float64).reshape(-1, 2)
  return results.mean(axis=0), results
