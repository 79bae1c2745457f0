"""Times the observation / legal-mask writer (expand_kernel) for the library named by BSG_LIB."""
import os, sys, torch
sys.path.insert(0, os.getcwd())
import open_spiel_1_b200 as bsg
from open_spiel_1_b200.roots import make_roots, ROOT_DEPTHS
dev = torch.device("cuda", 0)
def t(fn, reps=10):
  for _ in range(3): fn()
  torch.cuda.synchronize()
  a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
  a.record()
  for _ in range(reps): fn()
  b.record(); torch.cuda.synchronize()
  return a.elapsed_time(b) / reps
out = []
for game, n in (("breakthrough", 1 << 22), ("hex", 1 << 20), ("connect_four", 1 << 22)):
  eng = bsg.Engine(game, 0)
  st, _ = make_roots(eng, n, *ROOT_DEPTHS[game], seed=3)
  obs = torch.empty((n, eng.obs_size), dtype=torch.float32, device=dev)
  ms = t(lambda: eng.observations(st, n, None, out=obs))
  b = n * (eng.obs_size * 4 + eng.state_bytes)
  out.append("%s obs %.3f ms %.0f GB/s" % (game, ms, b / ms / 1e6))
  if game == "breakthrough":
    mask = torch.empty((n, eng.num_actions), dtype=torch.uint8, device=dev)
    ms = t(lambda: eng.observations(st, n, None, out=obs, legal_mask=mask))
    out.append("obs+mask %.3f ms %.0f GB/s" % (ms, (b + n * eng.num_actions) / ms / 1e6))
  eng.close()
print(os.environ.get("BSG_LIB", "default").split("/")[-1], " | ".join(out))
