"""Self-play actor with many games in flight on one GPU.

Mirrors `_play_game` of the reference's AlphaZero actor
(python/algorithms/alpha_zero/alpha_zero.py:196-228) and the learner's filing of its
trajectories (alpha_zero.py:366-369) for the evaluator of this engine (uniform priors, random
rollouts — `mcts.RandomRolloutEvaluator`, mcts.py:46-82):

    root   = bot.mcts_search(state)
    policy = explore_count of the root's children ** (1 / temperature), normalised
    action = root.best_child().action once `temperature_drop` moves were made, else ~ policy
    record = TrainInput(observation, legals_mask, policy, value = the game's final returns[0])

Game v is tree v of a GPU-resident forest (csrc/bsg_forest.cuh) with per-tree roots; one search
advances every running game by a move and the kernels of csrc/bsg_actor.cuh read the root
statistics, choose and apply the moves and log packed states / policies on the device.  At the
end the records are laid out game-major and the batched observation writer expands the logged
states straight into the replay buffer's tensors.  torch is used for device memory only.

`reference_style_game` is the same game played the reference's way — one search per move by a
host `MCTSBot` — with this module's (pinned) sampling rule; the parity test holds the device
actor to it record for record.
"""
import ctypes

import numpy as np

from . import _cabi as C
from .forest import _MAX_CHILDREN


def sample_action(visits_by_action, r, temperature=1.0):
  """The actor's sampling rule (csrc/bsg_actor.cuh): inverse CDF over ascending action ids with
  the 32-bit draw r; exact integer arithmetic at temperature 1."""
  if temperature == 1.0:
    total = int(sum(int(x) for x in visits_by_action))
    k = (int(r) * total) >> 32
    cum = 0
    for a, x in enumerate(visits_by_action):
      cum += int(x)
      if cum > k:
        return a
    raise ValueError("no visits")
  w = [float(x) ** (1.0 / temperature) for x in visits_by_action]
  wsum = 0.0
  for x in w:
    wsum += x
  u = ((float(r) + 0.5) * (1.0 / 4294967296.0)) * wsum
  cum, action = 0.0, -1
  for a, x in enumerate(w):
    cum += x
    if x > 0.0:
      action = a
    if cum > u:
      break
  return action


def reference_style_game(game, bot, temperature, temperature_drop, sample_seed, game_id):
  """alpha_zero.py:196-228 with a given bot: returns (records, returns); a record is
  (observation, current_player, legals_mask, action, policy float64 [A], root value)."""
  from .engine import draw
  records = []
  state = game.new_initial_state()
  move = 0
  while not state.is_terminal():
    root = bot.mcts_search(state)
    visits = np.zeros(game.num_distinct_actions())
    for c in root.children:
      visits[c.action] = c.explore_count
    policy = visits ** (1 / temperature)
    policy /= policy.sum()
    if move >= temperature_drop:
      action = root.best_child().action
    else:
      action = sample_action(visits.tolist(), draw(sample_seed, game_id, move), temperature)
    records.append((state.observation_tensor(), state.current_player(), state.legal_actions_mask(),
                    action, policy, root.total_reward / root.explore_count))
    state.apply_action(action)
    move += 1
  return records, state.returns()


class ForestActor:
  """n_games self-play games at a time, one forest tree each.

  Args:
    game: spiel.Game / pyspiel.Game.
    evaluator: GpuRolloutEvaluator — supplies the engine, n_rollouts, the Philox seed and the
      playout-id base (tree v draws from base + (v << 40), as forest.ForestMCTS).
    uct_c, max_simulations, solve, leaf_batch: the per-move search (MCTSBot arguments; the
      reference's actor runs with solve=False, alpha_zero.py:184-193).
    temperature, temperature_drop: alpha_zero.py Config fields.
    seed: tree v shuffles priors with seed + v; sample_seed: Philox key of the sampled moves.
  """

  def __init__(self, game, evaluator, n_games, uct_c=2.0, max_simulations=64, temperature=1.0,
               temperature_drop=10, solve=False, seed=0, sample_seed=0, leaf_batch=8, max_moves=None):
    import torch
    self.engine = evaluator.engine
    self.game = game
    self.n_games = int(n_games)
    self.sims = int(max_simulations)
    self.temperature = float(temperature)
    self.temperature_drop = int(temperature_drop)
    self.sample_seed = int(sample_seed)
    self.num_actions = game.num_distinct_actions()
    self.max_moves = int(max_moves or self.engine.max_len)
    self.games_played = 0
    e = self.engine
    cap = (self.sims + 2) * _MAX_CHILDREN[e.gid] + 8
    h = C.vp()
    C.check(C.lib.bsg_forest_create(e._h, self.n_games, cap, self.sims + 8, float(uct_c),
                                    int(bool(solve)), int(evaluator.n_rollouts), int(leaf_batch),
                                    int(evaluator.seed), int(seed), int(evaluator.next_playout),
                                    ctypes.byref(h)), e._h)
    self._h = h
    dev, T, M, A = e.torch_device, self.n_games, self.max_moves, self.num_actions
    z = lambda shape, dt: torch.zeros(shape, dtype=dt, device=dev)
    self.states = z(T * e.state_bytes, torch.uint8)
    self.alive = z(T, torch.uint8)
    self.length = z(T, torch.int32)
    self.outcome = z(T, torch.int32)
    self.n_alive = z(1, torch.int32)
    self.state_log = z(M * T * e.state_bytes, torch.uint8)
    self.policy_log = z((M * T, A), torch.float32)
    self.est_log = z(M * T, torch.float32)
    self.action_log = z(M * T, torch.int32)
    self.weights = z((T, A), torch.float64)
    from .engine import pack_histories, interleave_packed
    one = pack_histories(e.gid, [[]])
    self._initial = e.to_device(interleave_packed(e.gid, [one] * T))

  def close(self):
    if getattr(self, "_h", None):
      C.lib.bsg_forest_destroy(self._h)
      self._h = None

  def __del__(self):
    try:
      self.close()
    except Exception:
      pass

  def _check(self, rc):
    if rc < 0:
      raise RuntimeError("bsg: " + C.lib.bsg_forest_last_error(self._h).decode())
    return rc

  def play_batch(self, buffer=None, want_extras=False):
    """Plays n_games games to the end.  Returns a dict of DEVICE tensors, records game-major
    (game 0's moves, then game 1's, ...): observation [R, C*H*W] f32, legals_mask [R, A] bool,
    policy [R, A] f32, value [R, 1] f32 (the game's final returns[0]); lengths / returns per
    game; with want_extras also action, player, root value estimate per record.  With `buffer`
    (replay.DeviceReplayBuffer) the records are appended to it."""
    import torch
    e = self.engine
    dev, T, M, A = e.torch_device, self.n_games, self.max_moves, self.num_actions
    stream = torch.cuda.ExternalStream(C.lib.bsg_engine_stream(e._h), device=dev)
    torch.cuda.current_stream(dev).synchronize()
    with torch.cuda.stream(stream):
      self.states.copy_(self._initial)
      self.alive.fill_(1)
      self.length.zero_()
      self.outcome.zero_()
    io = C.ActorIO()
    p = lambda t: C.vp(t.data_ptr())
    io.states, io.alive, io.length, io.outcome = p(self.states), p(self.alive), p(self.length), p(self.outcome)
    io.n_alive_dev, io.state_log, io.policy_log = p(self.n_alive), p(self.state_log), p(self.policy_log)
    io.est_log, io.action_log, io.weights = p(self.est_log), p(self.action_log), p(self.weights)
    io.n_games, io.max_moves, io.temperature_drop = T, M, self.temperature_drop
    io.temperature, io.sample_seed, io.game_id_base = self.temperature, self.sample_seed, self.games_played
    n_alive = ctypes.c_uint32(T)
    move = 0
    while n_alive.value > 0:
      if move >= M:
        raise RuntimeError("bsg: a game outlasted max_moves = %d" % M)
      self._check(C.lib.bsg_forest_search_roots(self._h, p(self.states), p(self.alive), self.sims))
      io.move = move
      self._check(C.lib.bsg_forest_actor_step(self._h, ctypes.byref(io), ctypes.byref(n_alive)))
      move += 1
    stream.synchronize()
    lengths = self.length.cpu().numpy().astype(np.int64)
    offsets_np = np.concatenate([[0], np.cumsum(lengths)[:-1]]).astype(np.int64)
    total = int(lengths.sum())
    self.games_played += T
    out_states = torch.zeros(total * e.state_bytes, dtype=torch.uint8, device=dev)
    offsets = torch.from_numpy(offsets_np).to(dev)
    # write straight into the replay buffer when the rows are contiguous there
    direct = (buffer is not None and buffer._head + total <= buffer.max_size and total <= buffer.max_size and
              (buffer._head * e.obs_size * 4) % 16 == 0)     # the batched writer stores 16 bytes at a time
    if direct:
      rows = slice(buffer._head, buffer._head + total)
      obs, policy, value = buffer.observation[rows], buffer.policy[rows], buffer.value[rows]
    else:
      obs = torch.empty((total, e.obs_size), dtype=torch.float32, device=dev)
      policy = torch.empty((total, A), dtype=torch.float32, device=dev)
      value = torch.empty((total, 1), dtype=torch.float32, device=dev)
    mask8 = torch.empty((total, A), dtype=torch.uint8, device=dev)
    est = torch.empty(total, dtype=torch.float32, device=dev) if want_extras else None
    action = torch.empty(total, dtype=torch.int32, device=dev) if want_extras else None
    player = torch.empty(total, dtype=torch.int32, device=dev) if want_extras else None
    em = C.ActorEmitIO()
    em.state_log, em.policy_log, em.est_log, em.action_log = p(self.state_log), p(self.policy_log), p(self.est_log), p(self.action_log)
    em.length, em.outcome, em.offsets = p(self.length), p(self.outcome), p(offsets)
    em.out_states, em.out_policy, em.out_value = p(out_states), p(policy), p(value)
    em.out_est = p(est) if want_extras else None
    em.out_action = p(action) if want_extras else None
    em.out_player = p(player) if want_extras else None
    em.n_games, em.max_moves, em.total = T, M, total    # M: the capacity the logs were written with
    torch.cuda.current_stream(dev).synchronize()      # the tensors above were made on torch's stream
    C.check(C.lib.bsg_actor_emit(e._h, ctypes.byref(em)), e._h)
    C.check(C.lib.bsg_observations(e._h, p(out_states), total, None, p(obs), p(mask8),
                                   C.vp(C.lib.bsg_engine_stream(e._h))), e._h)
    stream.synchronize()
    mask = mask8.to(torch.bool)
    if buffer is not None:
      if direct:
        buffer.legals_mask[rows] = mask
        buffer.total_seen += total
        buffer._head = (buffer._head + total) % buffer.max_size
        buffer._size = min(buffer.max_size, buffer._size + total)
      else:
        buffer.extend(obs, mask, policy, value)
    out = {"observation": obs, "legals_mask": mask, "policy": policy, "value": value, "total": total,
           "lengths": lengths, "offsets": offsets_np, "returns": self.outcome.cpu().numpy().copy(),
           "moves": move}
    if want_extras:
      out.update({"action": action, "player": player, "estimate": est, "states": out_states})
    return out
