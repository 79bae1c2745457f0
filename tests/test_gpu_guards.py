"""Out-of-bounds guard tests (compute-sanitizer is closed on this pool: profiles/
r2q_sanitizer_unavailable.txt).  Every caller-owned output of the kernels is a slice of a larger
buffer filled with a canary byte; after the launch the bytes before and after the slice must be
untouched.  Sizes are ragged on purpose (not multiples of the 256-state chunk, of a warp, or of
16 bytes) — the three-buffer bulk-store ring of the observation writer and the tail code of its
last chunk are what an overrun would come from."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu
GAMES = ["tic_tac_toe", "connect_four", "breakthrough", "hex", "pentago", "othello", "y", "havannah", "go(board_size=9)"]
CANARY = 0xA5
PAD = 4096


class Guarded:
  """A device tensor of `numel` elements of `dtype` inside a canary-filled buffer, 16-byte aligned."""

  def __init__(self, torch, dev, numel, dtype):
    self.torch = torch
    self.item = torch.empty(0, dtype=dtype).element_size()
    self.nbytes = int(numel) * self.item
    self.raw = torch.full((PAD + self.nbytes + PAD,), CANARY, dtype=torch.uint8, device=dev)
    self.t = self.raw[PAD:PAD + self.nbytes].view(dtype)

  def intact(self):
    return bool((self.raw[:PAD] == CANARY).all()) and bool((self.raw[PAD + self.nbytes:] == CANARY).all())


@pytest.mark.parametrize("game", GAMES)
@pytest.mark.parametrize("n", [1, 255, 257, 1000, 4099])
def test_writers_stay_inside_their_buffers(game, n):
  import torch
  import open_spiel_1_b200 as bsg
  from open_spiel_1_b200.roots import make_roots, ROOT_DEPTHS
  eng = bsg.Engine(game, 0)
  dev = eng.torch_device
  dmin, dmax = ROOT_DEPTHS[game]
  states, _ = make_roots(eng, n, dmin, dmax, seed=n)
  # observation planes + legal mask
  obs = Guarded(torch, dev, n * eng.obs_size, torch.float32)
  mask = Guarded(torch, dev, n * eng.num_actions, torch.uint8)
  players = torch.from_numpy((np.arange(n) % 2).astype(np.uint8)).to(dev)
  eng.observations(states, n, players, out=obs.t.view((n,) + tuple(eng.obs_shape)),
                   legal_mask=mask.t.view(n, eng.num_actions))
  torch.cuda.synchronize()
  assert obs.intact() and mask.intact()
  assert set(np.unique(obs.t.cpu().numpy()).tolist()) <= {0.0, 1.0}
  assert set(np.unique(mask.t.cpu().numpy()).tolist()) <= {0, 1}
  # playout outputs, lean and full kernels
  R = 3
  stride = 7                                       # shorter than most playouts: the log must clip
  ret0 = Guarded(torch, dev, n * R, torch.int8)
  length = Guarded(torch, dev, n * R, torch.uint8)
  term = Guarded(torch, dev, n * R, torch.uint8)
  acts = Guarded(torch, dev, n * R * stride, torch.int16)
  rs = Guarded(torch, dev, n, torch.int32)
  rl = Guarded(torch, dev, n, torch.int32)
  tot = Guarded(torch, dev, 4, torch.int64)
  fin = Guarded(torch, dev, n * R * eng.state_bytes, torch.uint8)
  for g_ in (rs, rl, tot):
    g_.t.zero_()
  eng.playouts(states, n, R, seed=3, first_playout=9, ret0=ret0.t, length=length.t, totals=tot.t)
  eng.playouts(states, n, R, seed=3, first_playout=9, ret0=ret0.t, length=length.t, terminal=term.t,
               actions=acts.t.view(n * R, stride), root_sum_ret0=rs.t, root_sum_steps=rl.t,
               final_states=fin.t)
  torch.cuda.synchronize()
  for g_ in (ret0, length, term, acts, rs, rl, tot, fin):
    assert g_.intact()
  assert int(tot.t[0]) == n * R          # only the first launch was given totals
  eng.close()


@pytest.mark.parametrize("game", ["connect_four", "hex", "breakthrough"])
def test_trajectory_compaction_and_actor_logs_stay_inside(game):
  import ctypes
  import torch
  import open_spiel_1_b200 as bsg
  from open_spiel_1_b200 import _cabi as C
  from open_spiel_1_b200.roots import make_roots, ROOT_DEPTHS
  eng = bsg.Engine(game, 0)
  dev = eng.torch_device
  n = 777
  dmin, dmax = ROOT_DEPTHS[game]
  states, _ = make_roots(eng, n, dmin, dmax, seed=5)
  stride = eng.max_len
  length = Guarded(torch, dev, n, torch.uint8)
  strided = Guarded(torch, dev, n * stride * eng.state_bytes, torch.uint8)
  acts = Guarded(torch, dev, n * stride, torch.int16)
  off = Guarded(torch, dev, n + 1, torch.int64)
  eng.playouts(states, n, 1, seed=8, length=length.t, state_log=strided.t, state_log_stride=stride,
               actions=acts.t.view(n, stride))
  total_c = ctypes.c_int64(0)
  C.check(C.lib.bsg_trajectory_scan(eng._h, eng._p(length.t), n, eng._p(off.t), ctypes.byref(total_c),
                                    eng._stream_ptr(None)), eng._h)
  total = int(total_c.value)
  lens = length.t.cpu().numpy().astype(np.int64)
  assert total == int(lens.sum())
  assert np.array_equal(off.t.cpu().numpy(), np.concatenate([[0], np.cumsum(lens)]))
  out_states = Guarded(torch, dev, total * eng.state_bytes, torch.uint8)
  out_a = Guarded(torch, dev, total, torch.int64)
  out_p = Guarded(torch, dev, total, torch.int64)
  out_s = Guarded(torch, dev, total, torch.int64)
  C.check(C.lib.bsg_trajectory_compact(eng._h, eng._p(strided.t), n, stride, eng._p(off.t), total,
                                       eng._p(acts.t), stride, eng._p(out_states.t), eng._p(out_a.t),
                                       eng._p(out_p.t), eng._p(out_s.t), eng._stream_ptr(None)), eng._h)
  torch.cuda.synchronize()
  for g_ in (length, strided, acts, off, out_states, out_a, out_p, out_s):
    assert g_.intact()
  assert np.array_equal(out_p.t.cpu().numpy(), np.repeat(np.arange(n), lens))
  eng.close()
