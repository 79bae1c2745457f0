#!/bin/bash
# A/B of library builds under scratch/v on one game
G=${1:-y}; P=${2:-524288}
for f in scratch/v/*.so; do
  BSG_LIB=$PWD/$f python bench.py --game $G --playouts $P --no-cpu-baseline --no-extras --steps 10 --warmup 3 2>&1 | python -c "
import sys, json
for l in sys.stdin:
  if l.startswith('{'):
    d = json.loads(l); print('$f', '%.4g steps/s %.4f ms' % (d['value'], d['ms_per_step']))
"
done
