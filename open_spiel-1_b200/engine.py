"""Engine: the device side of the C ABI, driven from Python.

PyTorch is used only for device memory, streams and handing observation tensors to models;
all game work happens in the CUDA kernels of libbsg.so.  Nothing here has a CPU fallback:
constructing an Engine without a CUDA device raises RuntimeError.
"""
import ctypes

import numpy as np

from . import _cabi as C

GAME_NAMES = ("tic_tac_toe", "connect_four", "breakthrough", "hex", "pentago", "othello", "y", "havannah", "go(board_size=9)")


def game_id(game):
  """Accepts a game id, a reference game string ("hex(board_size=11)") or a Game-like object."""
  if isinstance(game, int):
    return game
  if not isinstance(game, str):
    game = str(game)
  gid = C.lib.bsg_game_id(game.encode())
  if gid < 0:
    raise RuntimeError("bsg: unsupported game '%s' (supported: tic_tac_toe, connect_four, "
                       "breakthrough 8x8, hex 11x11, pentago, othello, y 19, havannah 8, go 9x9)" % game)
  return gid


def state_bytes(gid):
  return C.check(C.lib.bsg_state_bytes(gid))


_PLANES = {}


def plane_bytes(gid):
  """Element sizes of the packed SoA planes of a game (bsg_plane_bytes)."""
  if gid not in _PLANES:
    out = (ctypes.c_int * 16)()
    n = C.check(C.lib.bsg_plane_bytes(gid, out, 16))
    _PLANES[gid] = [int(out[i]) for i in range(n)]
  return _PLANES[gid]


def interleave_packed(gid, packs):
  """Single packed states (capacity 1 each) -> one packed SoA buffer of capacity len(packs)."""
  n = len(packs)
  sb = state_bytes(gid)
  rows = np.stack([np.ascontiguousarray(p, np.uint8).reshape(sb) for p in packs]) if n else np.zeros((0, sb), np.uint8)
  parts, off = [], 0
  for pb in plane_bytes(gid):
    parts.append(rows[:, off:off + pb].reshape(-1))
    off += pb
  return np.concatenate(parts) if parts else np.zeros(0, np.uint8)


def num_distinct_actions(gid):
  return C.check(C.lib.bsg_num_distinct_actions(gid))


def max_game_length(gid):
  return C.check(C.lib.bsg_max_game_length(gid))


def observation_shape(gid):
  shape = (ctypes.c_int * 3)()
  C.check(C.lib.bsg_observation_shape(gid, shape))
  return [shape[0], shape[1], shape[2]]


def observation_size(gid):
  return C.check(C.lib.bsg_observation_size(gid))


def pack_histories(game, histories):
  """Action histories (State.history()) -> packed SoA states, a uint8 numpy buffer of
  len(histories) * state_bytes.  Raises RuntimeError on an illegal action."""
  gid = game_id(game)
  n = len(histories)
  off = np.zeros(n + 1, np.int64)
  for i, h in enumerate(histories):
    off[i + 1] = off[i] + len(h)
  flat = np.zeros(max(int(off[-1]), 1), np.int64)
  for i, h in enumerate(histories):
    if len(h):
      flat[off[i]:off[i + 1]] = h
  out = np.zeros(max(n, 1) * state_bytes(gid), np.uint8)
  C.check(C.lib.bsg_pack_histories(gid, flat.ctypes.data, off.ctypes.data, n, out.ctypes.data, n))
  return out[:n * state_bytes(gid)]


def draw(seed, playout, t):
  """The t-th replay draw r(seed, playout, t) of the engine's Philox contract."""
  return int(C.lib.bsg_draw(seed, playout, t))


class HostBuffer:
  """Page-locked host memory from bsg_host_alloc as a numpy uint8 array (`.array`); freed by
  close() / garbage collection.  write_combined=True: fill it with ONE sequential write
  (`buf.array[:] = data`) and do not read it back."""

  def __init__(self, nbytes, write_combined=False):
    self.nbytes = int(nbytes)
    self.ptr = C.lib.bsg_host_alloc(self.nbytes, 1 if write_combined else 0)
    if not self.ptr:
      raise RuntimeError("bsg: cudaHostAlloc of %d bytes failed" % self.nbytes)
    self.array = np.ctypeslib.as_array((ctypes.c_uint8 * self.nbytes).from_address(self.ptr))

  def close(self):
    if getattr(self, "ptr", None):
      self.array = None
      C.lib.bsg_host_free(self.ptr)
      self.ptr = None

  def __del__(self):
    try:
      self.close()
    except Exception:  # pragma: no cover
      pass


def _torch():
  import torch  # local: only device paths need it
  return torch


class Engine:
  """One engine per (thread, GPU) — the threading contract of the reference evaluators."""

  def __init__(self, game, device=0):
    self.gid = game_id(game)
    if self.gid not in C.DEVICE_GAMES:
      raise RuntimeError("bsg: game id %d has host helpers only" % self.gid)
    torch = _torch()
    if not torch.cuda.is_available():
      raise RuntimeError("bsg: no CUDA device — the playout engine has no CPU fallback")
    self.device = int(device)
    self.torch_device = torch.device("cuda", self.device)
    handle = C.vp()
    rc = C.lib.bsg_create(self.gid, self.device, ctypes.byref(handle))
    self._h = handle
    if rc != C.OK:
      try:
        C.check(rc, handle if handle.value else None)
      finally:
        if handle.value:
          C.lib.bsg_destroy(handle)
        self._h = None
    self.state_bytes = state_bytes(self.gid)
    self.num_actions = num_distinct_actions(self.gid)
    self.obs_shape = observation_shape(self.gid)
    self.obs_size = observation_size(self.gid)
    self.max_len = max_game_length(self.gid)
    self.sm_count = C.lib.bsg_engine_sm_count(self._h)
    self.launches = 0  # kernels launched through this engine (bench.py's gpu_launches)

  def close(self):
    if getattr(self, "_h", None):
      C.lib.bsg_destroy(self._h)
      self._h = None

  def __del__(self):
    try:
      self.close()
    except Exception:  # pragma: no cover
      pass

  # ---- helpers
  def to_device(self, packed):
    """Packed host buffer (numpy uint8) -> device tensor."""
    torch = _torch()
    t = torch.from_numpy(np.ascontiguousarray(packed))
    return t.to(self.torch_device, non_blocking=False)

  def _stream_ptr(self, stream):
    torch = _torch()
    if stream is None:
      stream = torch.cuda.current_stream(self.torch_device)
    return C.vp(stream.cuda_stream)

  @staticmethod
  def _p(t):
    return None if t is None else C.vp(t.data_ptr())

  def set_value_only(self, on=True):
    """hex / y: the host-buffer playouts (playouts_host, hence GpuRolloutEvaluator) fill the
    board and read the winner off it — same returns, about a third of the instructions."""
    C.check(C.lib.bsg_engine_set_value_only(self._h, 1 if on else 0), self._h)

  # ---- the hot path
  def playouts(self, roots, n_roots, rollouts=1, seed=0, first_playout=0, *, ret0=None,
               length=None, terminal=None, actions=None, root_sum_ret0=None,
               root_sum_steps=None, totals=None, max_steps=0, step_limit=None, r_table=None,
               final_states=None, final_before_terminal=False, state_log=None,
               state_log_off=None, root_playout_base=None, value_only=False, state_log_stride=0,
               stream=None):
    """Launches n_roots * rollouts random playouts (bsg_playouts).  All tensors are device
    tensors owned by the caller; see include/bsg.h for their meaning.  Asynchronous."""
    io = C.PlayoutIO()
    io.max_steps = int(max_steps)
    io.step_limit = self._p(step_limit)
    io.r_table = self._p(r_table)
    io.r_stride = int(r_table.shape[1]) if r_table is not None else 0
    io.ret0 = self._p(ret0)
    io.length = self._p(length)
    io.terminal = self._p(terminal)
    io.actions = self._p(actions)
    io.actions_stride = int(actions.shape[1]) if actions is not None else 0
    io.root_sum_ret0 = self._p(root_sum_ret0)
    io.root_sum_steps = self._p(root_sum_steps)
    io.final_states = self._p(final_states)
    io.final_before_terminal = 1 if final_before_terminal else 0
    io.totals = self._p(totals)
    io.state_log = self._p(state_log)
    io.state_log_cap = int(state_log.numel() // self.state_bytes) if state_log is not None else 0
    io.state_log_off = self._p(state_log_off)
    io.root_playout_base = self._p(root_playout_base)
    io.value_only = 1 if value_only else 0   # hex / y: fill the board, read the winner off it
    io.state_log_stride = int(state_log_stride)
    C.check(C.lib.bsg_playouts(self._h, self._p(roots), int(n_roots), int(rollouts),
                               int(seed) & (2**64 - 1), int(first_playout), ctypes.byref(io),
                               self._stream_ptr(stream)), self._h)
    self.launches += 1

  def noalloc_max_playouts(self):
    """Playouts one playouts_value_only_noalloc call can run (the static scratch's capacity)."""
    return C.check(C.lib.bsg_noalloc_max_playouts(self.gid))

  def playouts_value_only_noalloc(self, roots, n_roots, rollouts=1, seed=0, first_playout=0, *,
                                  ret0=None, length=None, root_sum_ret0=None, root_sum_steps=None,
                                  totals=None, stream=None):
    """bsg_playouts_value_only_noalloc: value-only playouts (hex, y) whose scratch is a static
    device array of the library — the call allocates nothing.  Same outputs as
    playouts(..., value_only=True); at most noalloc_max_playouts() playouts per call; calls
    that overlap on the GPU race on the shared array (include/bsg.h)."""
    C.check(C.lib.bsg_playouts_value_only_noalloc(
        self._h, self._p(roots), int(n_roots), int(rollouts), int(seed) & (2**64 - 1),
        int(first_playout), self._p(ret0), self._p(length), self._p(root_sum_ret0),
        self._p(root_sum_steps), self._p(totals), self._stream_ptr(stream)), self._h)
    self.launches += 2

  def observations(self, states, n, players=None, out=None, legal_mask=None, want_mask=False,
                   stream=None):
    """Batched ObservationTensor (+ LegalActionsMask) writer.  Returns (obs, mask) device
    tensors of shape [n, C, H, W] float32 and [n, NumDistinctActions] uint8."""
    torch = _torch()
    if out is None:
      out = torch.empty((n,) + tuple(self.obs_shape), dtype=torch.float32,
                        device=self.torch_device)
    if want_mask and legal_mask is None:
      legal_mask = torch.empty((n, self.num_actions), dtype=torch.uint8,
                               device=self.torch_device)
    C.check(C.lib.bsg_observations(self._h, self._p(states), int(n), self._p(players),
                                   self._p(out), self._p(legal_mask), self._stream_ptr(stream)),
            self._h)
    self.launches += 1 + (1 if legal_mask is not None else 0)
    return out, legal_mask

  def _traj_buffer(self, store, name, numel, dtype):
    """A cached device tensor of at least `numel` elements (grown, never shrunk)."""
    torch = _torch()
    t = store.get(name)
    if t is None or t.numel() < numel or t.dtype != dtype:
      t = torch.empty(max(int(numel), 1), dtype=dtype, device=self.torch_device)
      store[name] = t
    return t

  def trajectories(self, roots, n_roots, rollouts=1, seed=0, first_playout=0, want_mask=True,
                   want_actions=True, out=None):
    """Self-play records of every decision point of n_roots * rollouts random playouts, as
    the AlphaZero actor keeps them (python/algorithms/alpha_zero/alpha_zero.py:196-228):
    observation of the player to move, legal-action mask, the action taken and the final
    return for player 0.  Everything stays on the device and every playout is played ONCE:
    the launch logs the packed state before every move (and the action) at a fixed stride per
    playout, a scan of the lengths gives dense offsets, a compaction kernel gathers record
    off[p] + t, and the batched writer expands the compact log (bsg_trajectory_*; no PyTorch
    operator runs between the first launch and the last).

    `out`: a dict returned by an earlier call — its tensors are reused when they are large
    enough (a steady-state actor loop then allocates nothing).

    Returns a dict of device tensors: obs [T, C, H, W] f32, mask [T, A] u8 (or None),
    actions [T] int64 (or None), playout [T] / step [T] int64 (owner and move number of each
    record), ret0 [n] int8, length [n] uint8, offsets [n] int64, states (packed log, capacity T).
    The tensors are views of buffers the engine (or `out`) owns: copy what must outlive the
    next call."""
    torch = _torch()
    dev = self.torch_device
    n = int(n_roots) * int(rollouts)
    stride = 120 if self.gid == C.OTHELLO else int(self.max_len)   # othello: placements + passes
    sc = self.__dict__.setdefault("_traj_scratch", {})
    ret0 = self._traj_buffer(sc, "ret0", n, torch.int8)[:n]
    length = self._traj_buffer(sc, "length", n, torch.uint8)[:n]
    strided = self._traj_buffer(sc, "strided", n * stride * self.state_bytes, torch.uint8)[:n * stride * self.state_bytes]
    acts = self._traj_buffer(sc, "acts", n * stride, torch.int16)[:n * stride].view(n, stride) if want_actions else None
    off = self._traj_buffer(sc, "off", n + 1, torch.int64)[:n + 1]
    stream = self._stream_ptr(None)
    self.playouts(roots, n_roots, rollouts, seed, first_playout, ret0=ret0, length=length,
                  state_log=strided, state_log_stride=stride, actions=acts)
    total_c = ctypes.c_int64(0)
    C.check(C.lib.bsg_trajectory_scan(self._h, self._p(length), n, self._p(off), ctypes.byref(total_c),
                                      stream), self._h)
    total = int(total_c.value)
    store = out if out is not None else self.__dict__.setdefault("_traj_out", {})
    store = store.setdefault("_buffers", {}) if out is not None else store
    t1 = max(total, 1)
    states = self._traj_buffer(store, "states", t1 * self.state_bytes, torch.uint8)[:t1 * self.state_bytes]
    obs = self._traj_buffer(store, "obs", t1 * self.obs_size, torch.float32)[:t1 * self.obs_size].view(
        (t1,) + tuple(self.obs_shape))
    mask = self._traj_buffer(store, "mask", t1 * self.num_actions, torch.uint8)[:t1 * self.num_actions].view(
        t1, self.num_actions) if want_mask else None
    actions = self._traj_buffer(store, "actions", t1, torch.int64)[:t1] if want_actions else None
    owner = self._traj_buffer(store, "playout", t1, torch.int64)[:t1]
    step = self._traj_buffer(store, "step", t1, torch.int64)[:t1]
    if total:
      C.check(C.lib.bsg_trajectory_compact(self._h, self._p(strided), n, stride, self._p(off), total,
                                           self._p(acts), stride if want_actions else 0,
                                           self._p(states), self._p(actions), self._p(owner),
                                           self._p(step), stream), self._h)
      self.observations(states, total, None, out=obs, legal_mask=mask)
      self.launches += 2
    res = {"obs": obs[:total] if total else None, "mask": mask[:total] if (want_mask and total) else None,
           "actions": actions[:total] if (want_actions and total) else None,
           "playout": owner[:total], "step": step[:total], "ret0": ret0, "length": length,
           "offsets": off[:n], "states": states, "total": total}
    if out is not None:
      res["_buffers"] = store
    return res

  # ---- host-buffer entry points (copies inside the call; synchronous)
  def playouts_host(self, packed, n_roots, rollouts=1, seed=0, first_playout=0,
                    want_root_sums=False):
    """bsg_playouts_host: packed numpy roots in, totals {playouts, steps, sum_ret0, stuck} out."""
    packed = np.ascontiguousarray(packed)
    totals = np.zeros(4, np.int64)
    rr = rs = None
    if want_root_sums:
      rr = np.zeros(n_roots, np.int32)
      rs = np.zeros(n_roots, np.int32)
    C.check(C.lib.bsg_playouts_host(self._h, packed.ctypes.data, int(n_roots), int(rollouts),
                                    int(seed) & (2**64 - 1), int(first_playout),
                                    totals.ctypes.data, None if rr is None else rr.ctypes.data,
                                    None if rs is None else rs.ctypes.data), self._h)
    self.launches += 1
    return totals, rr, rs

  def playouts_host_results(self, packed, n_roots, rollouts=1, seed=0, first_playout=0, ret0=None,
                            length=None):
    """bsg_playouts_host_results: packed numpy roots in; totals and the per-playout return for
    player 0 (int8) and move count (uint8) out, as numpy arrays of n_roots * rollouts entries.
    Pass page-locked arrays as `ret0` / `length` to have them written in place."""
    packed = np.ascontiguousarray(packed)
    n = int(n_roots) * int(rollouts)
    totals = np.zeros(4, np.int64)
    if ret0 is None:
      ret0 = np.empty(n, np.int8)
    if length is None:
      length = np.empty(n, np.uint8)
    C.check(C.lib.bsg_playouts_host_results(self._h, packed.ctypes.data, int(n_roots), int(rollouts),
                                            int(seed) & (2**64 - 1), int(first_playout),
                                            totals.ctypes.data, ret0.ctypes.data, length.ctypes.data),
            self._h)
    self.launches += 1
    return totals, ret0, length

  def compact_roots(self, packed, n_roots):
    """Packed roots -> the game's compact transfer format (connect_four: 49 bits per root, back
    to back)."""
    packed = np.ascontiguousarray(packed)
    out = np.zeros(int(C.check(C.lib.bsg_compact_buffer_bytes(self.gid, int(n_roots)))), np.uint8)
    C.check(C.lib.bsg_compact_roots(self.gid, packed.ctypes.data, int(n_roots), out.ctypes.data))
    return out

  def expand_roots(self, compact, n_roots):
    """The inverse of compact_roots, through the decode the kernel runs."""
    compact = np.ascontiguousarray(compact)
    out = np.zeros(int(n_roots) * self.state_bytes, np.uint8)
    C.check(C.lib.bsg_expand_roots(self.gid, compact.ctypes.data, int(n_roots), out.ctypes.data))
    return out

  def playouts_host_compact(self, compact, n_roots, rollouts=1, seed=0, first_playout=0, ret0=None,
                            length=None, want_results=False):
    """bsg_playouts_host_compact: compact numpy roots in; totals (and, with want_results or
    given arrays, the per-playout int8 return / uint8 move count) out."""
    compact = np.ascontiguousarray(compact)
    n = int(n_roots) * int(rollouts)
    totals = np.zeros(4, np.int64)
    if want_results and ret0 is None:
      ret0 = np.empty(n, np.int8)
    if want_results and length is None:
      length = np.empty(n, np.uint8)
    C.check(C.lib.bsg_playouts_host_compact(self._h, compact.ctypes.data, int(n_roots), int(rollouts),
                                            int(seed) & (2**64 - 1), int(first_playout), totals.ctypes.data,
                                            None if ret0 is None else ret0.ctypes.data,
                                            None if length is None else length.ctypes.data), self._h)
    self.launches += 1
    return totals, ret0, length

  def observations_host(self, packed, n, players=None, want_mask=False):
    packed = np.ascontiguousarray(packed)
    obs = np.empty((n,) + tuple(self.obs_shape), np.float32)
    mask = np.empty((n, self.num_actions), np.uint8) if want_mask else None
    pl = None if players is None else np.ascontiguousarray(players, np.uint8)
    C.check(C.lib.bsg_observations_host(self._h, packed.ctypes.data, int(n),
                                        None if pl is None else pl.ctypes.data, obs.ctypes.data,
                                        None if mask is None else mask.ctypes.data), self._h)
    self.launches += 1 + (1 if want_mask else 0)
    return obs, mask

  def int_peak(self, kind=2, iters=2000):
    """Measured integer issue rate (thread-instructions/s) for the roofline denominator."""
    out = ctypes.c_double()
    C.check(C.lib.bsg_int_peak(self._h, kind, iters, ctypes.byref(out)), self._h)
    return out.value
