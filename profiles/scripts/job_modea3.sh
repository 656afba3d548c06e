#!/bin/bash
for rep in 1 2; do
for lib in deep-rl-battlespace_b200/battlespace/libbattlespace.so ab/lib_branchy.so; do
  BATTLESPACE_LIB=$PWD/$lib python profiles/bench_configs.py modeA cfg5n 2>/dev/null | python -c "
import sys, json
for l in sys.stdin:
    d = json.loads(l)
    if 'envs' in d: print('$lib', d['envs'], d['l2_flushed_between_launches'], round(d['ms_step_kernel_median']*1e3,1), 'us', round(d['frac_algorithmic'],3))
    elif 'native' in d: print('$lib cfg5', round(d['native']['ms_per_tick']*1e3,2), 'us/tick')"
done
done
