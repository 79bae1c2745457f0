"""Kernel-only A/B timing: lean playout kernel per game on the bench's root recipe.
usage: [BSG_LIB=path.so] python scratch/kbench.py game[:playouts] ...   (CUDA events, 5 reps)"""
import sys, os, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import open_spiel_1_b200 as bsg
from open_spiel_1_b200.roots import make_roots, ROOT_DEPTHS
DEF = {"go(board_size=9)": 1 << 18, "havannah": 1 << 20, "connect_four": 1 << 24, "tic_tac_toe": 1 << 24, "breakthrough": 1 << 22, "hex": 1 << 20,
       "pentago": 1 << 22, "othello": 1 << 21, "y": 1 << 21}
out = {}
for spec in sys.argv[1:] or ["connect_four"]:
  g, _, n = spec.partition(":")
  n = int(n) if n else DEF[g]
  eng = bsg.Engine(g, 0)
  dmin, dmax = ROOT_DEPTHS[g]
  st, _ = make_roots(eng, n, dmin, dmax, seed=20240611 + (2 if g == "connect_four" else 4))  # bench.py seeds
  dev = eng.torch_device
  ret0 = torch.zeros(n, dtype=torch.int8, device=dev)
  length = torch.zeros(n, dtype=torch.uint8, device=dev)
  tot = torch.zeros(4, dtype=torch.int64, device=dev)
  f = lambda: eng.playouts(st, n, 1, seed=7, ret0=ret0, length=length, totals=tot)
  for _ in range(3): f()
  torch.cuda.synchronize(); tot.zero_()
  e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
  e0.record()
  for _ in range(5): f()
  e1.record(); torch.cuda.synchronize()
  ms = e0.elapsed_time(e1) / 5
  t = tot.cpu().numpy()
  out[spec] = {"ms": round(ms, 4), "gsteps_per_s": round(float(t[1]) / 5 / ms / 1e6, 2), "sum_ret": int(t[2])}
  eng.close()
print(os.environ.get("BSG_LIB", "default"), json.dumps(out))
