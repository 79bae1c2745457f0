import sys, os, time, json
sys.path.insert(0, '/root/repo')
import numpy as np, torch
import open_spiel_1_b200 as bsg
from open_spiel_1_b200.roots import make_roots
eng = bsg.Engine("connect_four", 0)
n = 1 << 24
st, _ = make_roots(eng, n, 8, 20, seed=20240613)
host = st.cpu().numpy()
comp_np = eng.compact_roots(host, n)
comp = torch.empty(comp_np.nbytes, dtype=torch.uint8, pin_memory=True); comp.numpy()[:] = comp_np
c = comp.numpy()
for i in range(3): eng.playouts_host_compact(c, n, 1, seed=1, first_playout=i * n)
torch.cuda.synchronize(); t0 = time.perf_counter(); gs = 0
for i in range(20):
  t, _, _ = eng.playouts_host_compact(c, n, 1, seed=1, first_playout=(3 + i) * n); gs += int(t[1])
torch.cuda.synchronize(); dt = time.perf_counter() - t0
print(os.environ.get("BSG_LIB", "default"), json.dumps({"e2e": gs / dt, "ms": dt / 20 * 1e3, "bytes": comp_np.nbytes}))
