"""bsg_playouts_value_only_noalloc: the value-only playouts (hex, y) with their scratch in a
static device array of the library, for callers that cannot call cudaMalloc.  Checked through
the C ABI against the existing entry point (bsg_playouts with io.value_only = 1, itself tied
to the exact kernels and the oracle in tests/test_gpu_parity.py): integer results, so the bar
is equality.  Two calls in a row share the array; the second one is smaller and uses other
playout ids, so stale boards of the first call would show.

This file sorts last on purpose.
"""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _run(eng, torch, fn, states, n, R, seed, first, **kw):
  dev = eng.torch_device
  out = dict(ret0=torch.zeros(n * R, dtype=torch.int8, device=dev),
             length=torch.zeros(n * R, dtype=torch.uint8, device=dev),
             root_sum_ret0=torch.zeros(n, dtype=torch.int32, device=dev),
             root_sum_steps=torch.zeros(n, dtype=torch.int32, device=dev),
             totals=torch.zeros(4, dtype=torch.int64, device=dev))
  fn(states, n, R, seed=seed, first_playout=first, **out, **kw)
  torch.cuda.synchronize(dev)
  return {k: v.cpu().numpy() for k, v in out.items()}


@pytest.mark.parametrize("game,n,R", [("hex", 40000, 4), ("y", 9000, 8)])
def test_noalloc_value_only_equals_the_allocating_entry_point(game, n, R):
  import torch
  import open_spiel_1_b200 as bsg
  from open_spiel_1_b200.roots import make_roots, ROOT_DEPTHS
  assert torch.cuda.is_available()
  eng = bsg.Engine(game, 0)
  cap = eng.noalloc_max_playouts()
  assert cap == {"hex": (16 << 20) // 16, "y": (16 << 20) // 48}[game]
  dmin, dmax = ROOT_DEPTHS[game]
  states, hist = make_roots(eng, n, dmin, dmax, seed=123, want_histories=True)
  # a few terminal roots in the batch (they park their board without playing)
  from oracle_lib import Oracle
  oracle = Oracle()
  for i in range(16):
    a, _, _ = oracle.replay(game, [], seed=4, playout=i)
    hist[n - 1 - i] = a
  states = eng.to_device(bsg.pack_histories(game, hist))
  # two calls in a row on the shared static array: a big one, then a smaller one with other ids
  for (m, seed, first) in ((n, 31, 1 << 33), (n // 3, 77, 5)):
    sub = eng.to_device(bsg.pack_histories(game, hist[n - m:]))
    want = _run(eng, torch, eng.playouts, sub, m, R, seed, first, value_only=True)
    got = _run(eng, torch, eng.playouts_value_only_noalloc, sub, m, R, seed, first)
    for k in want:
      assert np.array_equal(got[k], want[k]), (game, m, k)
    assert got["totals"][0] == m * R and got["totals"][3] == 0
    assert (got["ret0"] != 0).all()
  # totals-only call (what an evaluator asks for), outputs optional
  tot = torch.zeros(4, dtype=torch.int64, device=eng.torch_device)
  eng.playouts_value_only_noalloc(states, n, R, seed=31, first_playout=1 << 33, totals=tot)
  torch.cuda.synchronize()
  want = _run(eng, torch, eng.playouts, states, n, R, 31, 1 << 33, value_only=True)
  assert np.array_equal(tot.cpu().numpy(), want["totals"])
  # beyond the static scratch: refused with a clear error, nothing launched
  big_R = cap // n + 1
  with pytest.raises(RuntimeError, match="exceed the static scratch"):
    eng.playouts_value_only_noalloc(states, n, big_R, totals=tot)
  eng.close()


def test_noalloc_refuses_games_without_value_only_playouts():
  import torch
  import open_spiel_1_b200 as bsg
  eng = bsg.Engine("connect_four", 0)
  assert eng.noalloc_max_playouts() == 0
  root = eng.to_device(bsg.pack_histories("connect_four", [[]]))
  with pytest.raises(RuntimeError):
    eng.playouts_value_only_noalloc(root, 1, 4, totals=torch.zeros(4, dtype=torch.int64, device=eng.torch_device))
  eng.close()
