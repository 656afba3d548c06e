#!/bin/bash
L=deep-rl-battlespace_b200/battlespace/libbattlespace.so
bash profiles/scripts/job_ab.sh "$L ab/lib_pred.so"
bash profiles/scripts/job_ab.sh "ab/lib_b64.so" "BS_ROLLOUT_DYNSMEM=4352"
bash profiles/scripts/job_ab.sh "ab/lib_b64.so" "BS_ROLLOUT_DYNSMEM=0"
for lib in $L ab/lib_step16.so; do
  BATTLESPACE_LIB=$PWD/$lib python profiles/bench_configs.py modeA 2>/dev/null | python -c "
import sys, json
for l in sys.stdin:
    d = json.loads(l); print('$lib', d['envs'], d['l2_flushed_between_launches'], round(d['ms_step_kernel_median']*1e3,1), 'us', round(d['frac_algorithmic'],3))"
done
