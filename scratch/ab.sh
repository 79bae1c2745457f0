#!/bin/bash
# A/B: time the headline kernel for every library build under scratch/v/
for f in scratch/v/*.so; do
  BSG_LIB=$PWD/$f python bench.py --no-cpu-baseline --no-extras --steps 20 --warmup 5 2>&1 | python -c "
import sys, json
for l in sys.stdin:
  if l.startswith('{'):
    d = json.loads(l); print('$f', 'value %.4g  ms %.4f  frac %.3f' % (d['value'], d['ms_per_step'], d['roofline']['frac']))
  else: print(l[:200])
"
done
