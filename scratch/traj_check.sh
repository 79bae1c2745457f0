#!/bin/bash
# two bench runs; prints the trajectory write-out timing and the obs roofline of each
for i in 1 2; do python bench.py --no-cpu-baseline --steps 5 --warmup 3 2>&1 | tail -1 > gpurun_out/tr$i.json; done
python - <<'P'
import json
for i in (1, 2):
  d = json.loads(open("gpurun_out/tr%d.json" % i).read())
  print(d["trajectory_cfg3"]["ms"], d["trajectory_cfg3"]["records_per_s"], d["obs_roofline"]["frac"])
P
