#!/bin/bash
# Final evidence of a round: launch list of the bench command + --set full captures of the
# headline playout kernel and of the observation writer (each after a plain run exited 0).
TAG=${1:-r1m}
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-quality"
$CMD > gpurun_out/${TAG}_plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/${TAG}_plain.log; exit 1; }
ncu --metrics gpu__time_duration.sum --clock-control none -c 6000 --csv --log-file gpurun_out/${TAG}_launches.csv $CMD > gpurun_out/${TAG}_ncu_list.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:playout_kernel -s 4 -c 1 -o gpurun_out/${TAG}_c4 -f $CMD --no-extras > gpurun_out/${TAG}_ncu_c4.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:expand_kernel -s 2 -c 1 -o gpurun_out/${TAG}_obs -f $CMD > gpurun_out/${TAG}_ncu_obs.log 2>&1
ls -la gpurun_out | tail -12
