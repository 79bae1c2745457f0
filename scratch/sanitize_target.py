"""Target of the compute-sanitizer runs (profiles/r2*_sanitizer_*.txt): smoke() plus one forest
search with the engine-side merge, one self-play actor batch, one-pass trajectories and the
no-allocation entry point, all at small sizes."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import __graft_entry__ as g
g.smoke()
import open_spiel_1_b200 as bsg
from open_spiel_1_b200.forest import ForestMCTS
from open_spiel_1_b200.actor import ForestActor
for game in ("connect_four", "hex"):
  gm = bsg.load_game(game)
  eng = bsg.Engine(game, 0)
  ev = bsg.GpuRolloutEvaluator(eng, n_rollouts=8, seed=3)
  f = ForestMCTS(gm, 2.0, 64 * 32, ev, 64, sims_per_step=16, solve=True, seed=5, leaf_batch=4)
  root = f.mcts_search(gm.new_initial_state())
  assert sum(c.explore_count for c in root.children) > 0
  f.close()
  actor = ForestActor(gm, ev, 32, max_simulations=16, temperature_drop=4, leaf_batch=4)
  out = actor.play_batch(want_extras=True)
  assert out["total"] > 0
  actor.close()
  roots = eng.to_device(bsg.pack_histories(game, [[], [0], [1, 2]]))
  tr = eng.trajectories(roots, 3, 8, seed=4)
  assert tr["total"] > 0
  if game == "hex":
    tot = torch.zeros(4, dtype=torch.int64, device=eng.torch_device)
    eng.playouts_value_only_noalloc(roots, 3, 16, seed=2, totals=tot)
    torch.cuda.synchronize()
    assert int(tot[0]) == 48
  t, r, l = eng.playouts_host_results(bsg.pack_histories(game, [[], [0]]), 2, 4, seed=1)
  assert t[0] == 8
  eng.close()
print("sanitize target ok")
