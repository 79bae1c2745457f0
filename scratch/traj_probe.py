"""Where Engine.trajectories spends its time (cfg3 shape): CUDA events around each stage."""
import sys, os, json, ctypes
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import open_spiel_1_b200 as bsg
from open_spiel_1_b200 import _cabi as C
from open_spiel_1_b200.roots import make_roots
eng = bsg.Engine("breakthrough", 0)
n = 1 << 18
st, _ = make_roots(eng, n, 0, 40, seed=20240616)
for _ in range(3):
  tr = eng.trajectories(st, n, 1, seed=9)
torch.cuda.synchronize()
total = tr["total"]
dev = eng.torch_device
stride = eng.max_len
sc = eng._traj_scratch
ev = [torch.cuda.Event(enable_timing=True) for _ in range(8)]
names = []
def mark(i, name):
  ev[i].record(); names.append(name)
mark(0, "start")
eng.playouts(st, n, 1, 9, 0, ret0=sc["ret0"][:n], length=sc["length"][:n], state_log=sc["strided"], state_log_stride=stride,
             actions=sc["acts"][:n * stride].view(n, stride))
mark(1, "playouts(full, strided log)")
tc = ctypes.c_int64(0)
C.check(C.lib.bsg_trajectory_scan(eng._h, eng._p(sc["length"]), n, eng._p(sc["off"]), ctypes.byref(tc), eng._stream_ptr(None)))
mark(2, "scan")
out = eng._traj_out
C.check(C.lib.bsg_trajectory_compact(eng._h, eng._p(sc["strided"]), n, stride, eng._p(sc["off"]), total, eng._p(sc["acts"]), stride,
                                     eng._p(out["states"]), eng._p(out["actions"]), eng._p(out["playout"]), eng._p(out["step"]), eng._stream_ptr(None)))
mark(3, "compact")
obs = out["obs"][:total * eng.obs_size].view(total, 3, 8, 8)
mask = out["mask"][:total * eng.num_actions].view(total, eng.num_actions)
eng.observations(out["states"], total, None, out=obs)
mark(4, "obs planes")
eng.observations(out["states"], total, None, out=obs, legal_mask=mask)
mark(5, "obs planes + mask")
eng.playouts(st, n, 1, 9, 0, ret0=sc["ret0"][:n], length=sc["length"][:n])
mark(6, "playouts(lean)")
torch.cuda.synchronize()
res = {"records": total}
for i in range(1, 7):
  res[names[i]] = round(ev[i - 1].elapsed_time(ev[i]), 3)
print(json.dumps(res))
