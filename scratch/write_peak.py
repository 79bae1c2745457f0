"""Write-only HBM bandwidth ceiling: torch fill / zero / cudaMemset over 3.2 GB, and a copy."""
import torch
dev = torch.device("cuda", 0)
n = 3_221_225_472 // 4
x = torch.empty(n, dtype=torch.float32, device=dev)
y = torch.empty(n, dtype=torch.float32, device=dev)
def t(fn, reps=10):
  for _ in range(3): fn()
  torch.cuda.synchronize()
  a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
  a.record()
  for _ in range(reps): fn()
  b.record(); torch.cuda.synchronize()
  return a.elapsed_time(b) / reps
for name, fn, bytes_ in (("fill_(1.0)", lambda: x.fill_(1.0), n * 4), ("zero_()", lambda: x.zero_(), n * 4),
                         ("copy_", lambda: y.copy_(x), 2 * n * 4)):
  ms = t(fn)
  print("%-12s %.3f ms  %.1f GB/s" % (name, ms, bytes_ / ms / 1e6))
