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
  """Packed bitboards of one state.  The host mirror carries them (`packed()`), also when it
  was deserialised from a board string and has no history; a `pyspiel.State` crosses as its
  action history (pyspiel.cc:322) — except connect_four, whose Serialize() is the board itself
  (connect_four.cc:227) and therefore also right for a state the reference deserialised from
  a board string, where history() only holds the moves made since."""
  packed = getattr(state, "packed", None)
  if callable(packed):
    return packed()
  if gid == _engine.C.CONNECT_FOUR and callable(getattr(state, "serialize", None)):
    out = np.zeros(_engine.state_bytes(gid), np.uint8)
    _engine.C.check(_engine.C.lib.bsg_host_deserialize(gid, state.serialize().encode(),
                                                       out.ctypes.data, 1, 0))
    return out
  return _engine.pack_histories(gid, [list(state.history())])


def pack_states(gid, states):
  """Packed SoA buffer (capacity len(states)) of a list of states: one C call over the
  histories when no state carries its bitboards, the per-state planes interleaved otherwise."""
  if not any(callable(getattr(s, "packed", None)) for s in states) and gid != _engine.C.CONNECT_FOUR:
    return _engine.pack_histories(gid, [list(s.history()) for s in states])
  return _engine.interleave_packed(gid, [_packed_of(gid, s) for s in states])


class GpuRolloutEvaluator(Evaluator):
  """RandomRolloutEvaluator(n_rollouts, random_state) on the GPU.

  Args:
    game: game string / pyspiel.Game / spiel.Game, or an existing `engine.Engine`.
    n_rollouts: playouts per evaluate() call (the reference default is 1; the GPU pays one
      launch per call, so 256+ is the sensible regime — SURVEY.md §7.2).
    seed: Philox key.  Playout ids advance with every call, so results are reproducible for a
      given (seed, call sequence) and every playout is replayable by the oracle.
    device: CUDA device index.
    value_only: hex / y only — play plain stones until the board is full and read the winner
      off it (include/bsg.h, bsg_playout_io.value_only): evaluate() returns exactly the same
      values from the same playout ids; only `total_steps` counts moves to the full board.
  """

  def __init__(self, game, n_rollouts=1, seed=0, device=0, value_only=False):
    if isinstance(game, _engine.Engine):
      self.engine = game
    else:
      self.engine = _engine.Engine(game, device)
    if value_only:
      self.engine.set_value_only(True)
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
    packed = pack_states(self.engine.gid, states)
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


class AlphaZeroEvaluator(Evaluator):
  """`AlphaZeroEvaluator(game, model, cache_size)` of the reference
  (python/algorithms/alpha_zero/evaluator.py:38-83) over the batched GPU observation writer.

  The reference builds `state.observation_tensor()` and `state.legal_actions_mask()` as Python
  lists on the host, stacks a singleton batch and calls `model.inference(obs, mask)` which
  returns `(value [B, 1], policy [B, NumDistinctActions])` (alpha_zero/model.py).  Here the
  state crosses as its packed bitboards, the observation planes and the legal mask are written
  on the device by `bsg_observations` and handed to the model as DEVICE torch tensors
  `obs [B, C, H, W] float32`, `mask [B, A] bool` — no host copy of the planes.  `model` is any
  object with that `inference` method (returning torch tensors or numpy arrays).  Results are
  cached per packed state like the reference's LRU cache (evaluator.py:56-75).

  evaluate() returns `[value, -value]` exactly as the reference does (evaluator.py:77-80);
  prior() returns `[(action, policy[action])]` over the legal actions (evaluator.py:82-85).
  """

  def __init__(self, game, model, cache_size=2**16, device=0, engine=None):
    import collections
    if game.num_players() != 2:
      raise ValueError("Game must be for two players.")
    self.engine = engine if engine is not None else _engine.Engine(game, device)
    self._model = model
    self._cache = collections.OrderedDict()
    self._cache_size = int(cache_size)
    self.hits = self.misses = 0

  def clear_cache(self):
    self._cache.clear()

  def cache_info(self):
    return {"size": len(self._cache), "max_size": self._cache_size, "hits": self.hits,
            "misses": self.misses}

  def inference_batch(self, states):
    """(values [n], policies [n, A]) float64 for a list of states, one model call for all
    states that are not cached."""
    import torch
    gid = self.engine.gid
    packs = [np.ascontiguousarray(_packed_of(gid, s)) for s in states]
    keys = [p.tobytes() for p in packs]
    # this batch's answers are collected here first: inserting the misses may evict entries
    # (also this batch's own hits, or its own misses when the batch exceeds the cache)
    got = {}
    for k in keys:
      if k in self._cache and k not in got:
        got[k] = self._cache[k]
        self._cache.move_to_end(k)
    uniq = {}
    for i, k in enumerate(keys):
      if k not in got:
        uniq.setdefault(k, i)
    idx = list(uniq.values())
    if idx:
      n = len(idx)
      d_states = self.engine.to_device(_engine.interleave_packed(gid, [packs[i] for i in idx]))
      obs, mask = self.engine.observations(d_states, n, None, want_mask=True)
      value, policy = self._model.inference(obs, mask.to(torch.bool))
      value = torch.as_tensor(value).detach().to("cpu", torch.float64).reshape(n).numpy()
      policy = torch.as_tensor(policy).detach().to("cpu", torch.float64).reshape(n, -1).numpy()
      for j, i in enumerate(idx):
        got[keys[i]] = self._cache[keys[i]] = (float(value[j]), policy[j])
      while len(self._cache) > self._cache_size:
        self._cache.popitem(last=False)
    self.misses += len(idx)
    self.hits += len(states) - len(idx)
    vals = np.array([got[k][0] for k in keys], np.float64)
    pols = np.stack([got[k][1] for k in keys]) if keys else np.zeros((0, 0))
    return vals, pols

  def evaluate(self, state):
    v, _ = self.inference_batch([state])
    return np.array([v[0], -v[0]])

  def evaluate_batch(self, states):
    v, _ = self.inference_batch(states)
    return np.stack([v, -v], axis=1) if len(states) else np.zeros((0, 2))

  def prior(self, state):
    _, p = self.inference_batch([state])
    return [(action, p[0][action]) for action in state.legal_actions()]
