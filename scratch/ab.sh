#!/bin/bash
# A/B: time the headline kernel (and the other games) for every library build under scratch/v/
for f in scratch/v/*.so; do
  BSG_LIB=$PWD/$f python bench.py --no-cpu-baseline --steps 20 --warmup 5 2>&1 | python -c "
import sys, json
for l in sys.stdin:
  if l.startswith('{'):
    d = json.loads(l); og = d.get('other_games') or {}
    print('$f', 'c4 %.4g' % d['value'], ' '.join('%s %.4g' % (g, v['game_steps_per_s']) for g, v in og.items() if isinstance(v, dict)))
  else: print(l[:200])
"
done
