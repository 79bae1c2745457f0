#!/bin/bash
# quick GPU check: parity tests + short bench summary
python -m pytest tests -m gpu -x -q 2>&1 | tail -3
python bench.py --no-cpu-baseline "$@" 2>&1 | tee gpurun_out/bench_quick.log | python -c "
import sys, json
for l in sys.stdin:
  if l.startswith('{'):
    d = json.loads(l); print('value %.4g  ms %.4f  e2e %.4g  frac %.3f' % (d['value'], d['ms_per_step'], d['e2e']['value'], d['roofline']['frac']))
    for g, v in (d.get('other_games') or {}).items(): print(' ', g, v if isinstance(v, str) else '%.4g steps/s  %.3f ms  frac %.3f' % (v['game_steps_per_s'], v['ms'], v['int_roofline_frac']))
    if 'obs_roofline' in d: print('  obs frac %.3f' % d['obs_roofline']['frac'], d.get('trajectory_cfg3'))
  else: print(l[:300])
"
