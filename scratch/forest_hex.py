"""hex root-parallel MCTS on the GPU-resident forest, exact vs value-only leaves."""
import json, sys, time
import torch
import open_spiel_1_b200 as bsg
from open_spiel_1_b200.forest import ForestMCTS
for game in ("hex", "y"):
  g = bsg.load_game(game)
  state = g.new_initial_state()
  for vo in (False, True):
    eng = bsg.Engine(game, 0)
    ev = bsg.GpuRolloutEvaluator(eng, n_rollouts=64, seed=1, value_only=vo)
    trees, per_tree = 4096, 64
    f = ForestMCTS(g, 2.0, trees * per_tree, ev, trees, sims_per_step=32, solve=True, seed=3, leaf_batch=16,
                   node_capacity=(per_tree + 2) * (121 if game == "hex" else 190) + 8)
    f.mcts_search(state)                       # warm-up
    torch.cuda.synchronize(); t0 = time.time()
    root = f.mcts_search(state)
    torch.cuda.synchronize(); dt = time.time() - t0
    best = max(root.children, key=lambda c: c.explore_count).action
    print(json.dumps({"game": game, "value_only": vo, "trees": trees, "simulations": trees * per_tree,
                      "rollouts_per_leaf": 64, "seconds": dt, "simulations_per_s": trees * per_tree / dt,
                      "playouts_per_s": trees * per_tree * 64 / dt, "most_visited": best}))
    f.close(); eng.close()
