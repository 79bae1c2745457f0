#!/bin/bash
# ncu full capture of the playout kernel of a bench command (after a plain run)
TAG=${1:-prof}; shift
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-extras $@"
$CMD > gpurun_out/plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:playout_kernel -s 4 -c 1 -o gpurun_out/$TAG -f $CMD > gpurun_out/ncu_full.log 2>&1
tail -2 gpurun_out/ncu_full.log
