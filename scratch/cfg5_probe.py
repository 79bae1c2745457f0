"""cfg5 probe: one GPU holding `--trees` trees of 256 simulations (the per-GPU share of BASELINE
cfg5 at N = 4096 / trees GPUs); prints ms per move.  Run under ncu for the launch list."""
import argparse, json, sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import open_spiel_1_b200 as bsg
from open_spiel_1_b200.forest import ForestMCTS
ap = argparse.ArgumentParser()
ap.add_argument("--trees", type=int, default=512)
ap.add_argument("--sims", type=int, default=256)
ap.add_argument("--step", type=int, default=64)
ap.add_argument("--batch", type=int, default=8)
ap.add_argument("--moves", type=int, default=5)
a = ap.parse_args()
g = bsg.load_game("connect_four")
eng = bsg.Engine("connect_four", 0)
ev = bsg.GpuRolloutEvaluator(eng, n_rollouts=256, seed=3)
f = ForestMCTS(g, 2.0, a.trees * a.sims, ev, a.trees, sims_per_step=a.step, solve=True, seed=5, leaf_batch=a.batch)
st = g.new_initial_state()
f.mcts_search(st)
torch.cuda.synchronize()
t0 = time.perf_counter()
for _ in range(a.moves):
  f.mcts_search(st)
torch.cuda.synchronize()
dt = (time.perf_counter() - t0) / a.moves
print(json.dumps({"trees": a.trees, "sims_per_tree": a.sims, "batch": a.batch, "ms_per_move": dt * 1e3,
                  "sims_per_s": a.trees * a.sims / dt, "info": f.info()}))
