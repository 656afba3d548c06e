#!/bin/bash
# A/B of library variants on Mode A (bs_step at 65,536 / 1,048,576 envs) and the config-5 tick
# usage: job_modea3.sh "lib_a.so lib_b.so ..."
for rep in 1 2; do
for lib in $1; do
  BATTLESPACE_LIB=$PWD/$lib python profiles/bench_configs.py modeA cfg5n 2>/dev/null | python -c "
import sys, json
for l in sys.stdin:
    d = json.loads(l)
    if 'envs' in d: print('$lib', d['envs'], d['l2_flushed_between_launches'], round(d['ms_step_kernel_median']*1e3,1), 'us', round(d['frac_algorithmic'],3))
    elif 'native' in d: print('$lib cfg5', round(d['native']['ms_per_tick']*1e3,2), 'us/tick')"
done
done
