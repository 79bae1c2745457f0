#!/bin/bash
# cfg3 / cfg4 / othello / pentago weak scaling at 1, 2, 4, 8 GPUs and the GPU-resident forest (cfg5) at 8 GPUs
out=gpurun_out/r1k_scale.jsonl; : > $out
for spec in "breakthrough 4194304" "hex 1048576" "othello 2097152" "pentago 4194304"; do
  set -- $spec
  for n in 1 2 4 8; do
    if [ $n = 1 ]; then cmd="python bench.py"; else cmd="python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2950$n bench.py"; fi
    timeout 300 $cmd --gpus $n --steps 10 --warmup 3 --game $1 --playouts $2 --no-cpu-baseline --no-extras 2>gpurun_out/scale_err.log | grep '^{' | tail -1 >> $out
  done
done
for n in 8; do
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29611 bench_mcts.py --cfg 5 --forest 1 --trees 16384 --sims5 256 --leaf-batch 32 --sims-per-step 64 2>>gpurun_out/scale_err.log | grep '^{' | tail -1 > gpurun_out/r1k_mcts_forest_n8.json
done
python - <<'P'
import json
for l in open('gpurun_out/r1k_scale.jsonl'):
  d=json.loads(l); print(d['config'].get('game'), d['n_gpus'], '%.4g'%d['value'], '%.4f ms'%d['ms_per_step'])
print(open('gpurun_out/r1k_mcts_forest_n8.json').read()[:600])
P
