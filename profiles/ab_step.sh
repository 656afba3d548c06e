#!/bin/bash
# A/B of step-kernel build variants on the GPU box: cfg5 native loop + Mode A per library in ab/
for lib in "" $(ls ab/*.so 2>/dev/null); do
  BATTLESPACE_LIB=${lib:+$PWD/$lib} python profiles/bench_configs.py cfg5n modeA 2>&1 | python -c "
import sys,json
out=['${lib:-default}']
for l in sys.stdin:
    try: d=json.loads(l)
    except Exception: continue
    if 'native' in d: out.append('cfg5n %.4f' % d['native']['ms_per_tick'])
    else: out.append('modeA step %.4f team %.4f' % (d['ms_step_kernel'], d['ms_step_team_instinct_auto_reset']))
print(' | '.join(out))"
done
