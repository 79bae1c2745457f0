"""Drop-in evaluator for `open_spiel.python.algorithms.mcts.MCTSBot`.

Mirrors the reference interface (python/algorithms/mcts.py:29-82; C++ twin
algorithms/mcts.h:78-106):
    Evaluator.evaluate(state) -> array of returns, one per player
    Evaluator.prior(state)    -> mutable list of (action, probability) in legal-action order
`GpuRolloutEvaluator` is `RandomRolloutEvaluator`-compatible: evaluate() averages the terminal
returns of `n_rollouts` uniform-random playouts from `state` — all of them run in one launch of
the CUDA playout kernel — and prior() is uniform over `state.legal_actions()`.

`state` may be a real `pyspiel.State` or the host mirror `spiel.State`; it is never mutated
(the reference clones before rolling out, mcts.py:62).  It crosses the boundary as its action
history (pyspiel.cc:322) replayed into packed bitboards by the C ABI.
"""
import numpy as np

from . import engine as _engine


class Evaluator(object):
  """Abstract evaluator, same contract as mcts.Evaluator (mcts.py:29-43)."""

  def evaluate(self, state):
    raise NotImplementedError

  def prior(self, state):
    raise NotImplementedError


def _packed_of(gid, state):
  packed = getattr(state, "packed", None)
  if callable(packed):
    return packed()
  return _engine.pack_histories(gid, [list(state.history())])


class GpuRolloutEvaluator(Evaluator):
  """RandomRolloutEvaluator(n_rollouts, random_state) on the GPU.

  Args:
    game: game string / pyspiel.Game / spiel.Game, or an existing `engine.Engine`.
    n_rollouts: playouts per evaluate() call (the reference default is 1; the GPU pays one
      launch per call, so 256+ is the sensible regime — SURVEY.md §7.2).
    seed: Philox key.  Playout ids advance with every call, so results are reproducible for a
      given (seed, call sequence) and every playout is replayable by the oracle.
    device: CUDA device index.
  """

  def __init__(self, game, n_rollouts=1, seed=0, device=0):
    if isinstance(game, _engine.Engine):
      self.engine = game
    else:
      self.engine = _engine.Engine(game, device)
    self.n_rollouts = int(n_rollouts)
    if self.n_rollouts < 1:
      raise ValueError("n_rollouts must be >= 1")
    self.seed = int(seed)
    self.next_playout = 0      # global playout counter = Philox counter base
    self.last_playout_range = None
    self.total_steps = 0

  def evaluate(self, state):
    """Mean returns of n_rollouts random playouts from `state`, shape [2] float64."""
    packed = _packed_of(self.engine.gid, state)
    first = self.next_playout
    totals, _, _ = self.engine.playouts_host(packed, 1, self.n_rollouts, self.seed, first)
    self.next_playout += self.n_rollouts
    self.last_playout_range = (first, first + self.n_rollouts)
    self.total_steps += int(totals[1])
    v = float(totals[2]) / self.n_rollouts
    return np.array([v, -v])

  def evaluate_batch(self, states):
    """evaluate() for many leaves in one launch; returns [len(states), 2] float64.
    (The batched-leaf extension of SURVEY.md §8f row 1.)"""
    n = len(states)
    if n == 0:
      return np.zeros((0, 2))
    gid = self.engine.gid
    if all(callable(getattr(s, "packed", None)) for s in states) and n == 1:
      packed = states[0].packed()
    else:
      packed = _engine.pack_histories(gid, [list(s.history()) for s in states])
    first = self.next_playout
    totals, root_ret, _ = self.engine.playouts_host(packed, n, self.n_rollouts, self.seed, first,
                                                    want_root_sums=True)
    self.next_playout += n * self.n_rollouts
    self.last_playout_range = (first, first + n * self.n_rollouts)
    self.total_steps += int(totals[1])
    v = root_ret.astype(np.float64) / self.n_rollouts
    return np.stack([v, -v], axis=1)

  def prior(self, state):
    """Equal probability for every legal action, in ascending action order (mcts.py:76-82)."""
    legal_actions = state.legal_actions(state.current_player())
    return [(action, 1.0 / len(legal_actions)) for action in legal_actions]
