"""cfg3 trajectories on two engines and two streams: the second half's playout / log / compaction
kernels (ALU-bound, small) run under the first half's observation writer (HBM-bound).
Prints records/s and output GB/s of one call (the whole batch) and of the two-stream form."""
import sys, os, json, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import open_spiel_1_b200 as bsg
from open_spiel_1_b200.roots import make_roots
n = 1 << 18
parts = int(sys.argv[1]) if len(sys.argv) > 1 else 2
dev = torch.device("cuda:0")
e0 = bsg.Engine("breakthrough", 0)
st, _ = make_roots(e0, n, 0, 40, seed=20240616)
bytes_per_record = 4 * e0.obs_size + e0.num_actions + 8 + 8 + 8   # obs f32, mask u8, action, owner, step (int64)
for _ in range(3):
  tr = e0.trajectories(st, n, 1, seed=9)
torch.cuda.synchronize()
t0 = time.perf_counter()
for _ in range(5):
  tr = e0.trajectories(st, n, 1, seed=9)
torch.cuda.synchronize()
one = (time.perf_counter() - t0) / 5
total = tr["total"]
res = {"records": total, "one_call_ms": one * 1e3, "one_call_gbs": total * bytes_per_record / one / 1e9}
# the same playouts as `parts` calls of n / parts roots each, on alternating engines and streams
m = n // parts
engines = [e0, bsg.Engine("breakthrough", 0)]
streams = [torch.cuda.Stream(dev), torch.cuda.Stream(dev)]
roots = [make_roots(engines[i % 2], m, 0, 40, seed=20240616 + i)[0] for i in range(parts)]
outs = [{} for _ in range(parts)]
def run():
  tot = 0
  for i in range(parts):
    with torch.cuda.stream(streams[i % 2]):
      r = engines[i % 2].trajectories(roots[i], m, 1, seed=9, first_playout=i * m, out=outs[i])
      outs[i] = r
      tot += r["total"]
  return tot
for _ in range(3):
  run()
torch.cuda.synchronize()
t0 = time.perf_counter()
for _ in range(5):
  tot = run()
torch.cuda.synchronize()
two = (time.perf_counter() - t0) / 5
res.update({"parts": parts, "pipelined_records": tot, "pipelined_ms": two * 1e3, "pipelined_gbs": tot * bytes_per_record / two / 1e9})
print(json.dumps(res))
