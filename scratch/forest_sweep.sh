#!/bin/bash
# cfg5 with the GPU-resident forest: sweep trees / sims per tree / leaf batch
for cfg in "4096 1024 32 256" "16384 256 32 256" "32768 128 32 128" "65536 64 32 64"; do
  set -- $cfg
  python bench_mcts.py --cfg 5 --forest 1 --trees $1 --sims5 $2 --leaf-batch $3 --sims-per-step $4 2>&1 | tail -1 | tee -a gpurun_out/mcts_forest.jsonl | python -c "
import sys,json
l=sys.stdin.readline()
try:
  d=json.loads(l); print('$cfg', '%.4g sims/s %.4g playouts/s %.4g steps/s' % (d['simulations_per_s'], d['playouts_per_s'], d['game_steps_per_s']), '%.3f s' % d['seconds'], d['simulations'])
except Exception as e: print('$cfg', 'ERR', l[:300])"
done
