#!/bin/bash
python -m pytest tests/test_cuda_golden.py tests/test_dropin_api.py tests/test_edges.py tests/test_config1_full.py -m gpu -x -q 2>&1 | tail -2
L=deep-rl-battlespace_b200/battlespace/libbattlespace.so
for cfg in "$L BS_X=0" "$L BS_STEP_GENERAL=1" "ab/lib_lean14.so BS_X=0"; do
  set -- $cfg
  env $2 BATTLESPACE_LIB=$PWD/$1 python profiles/bench_configs.py modeA modeAloops 2>/dev/null | python -c "
import sys, json
for l in sys.stdin:
    d = json.loads(l)
    if 'envs' in d: print('$1 $2', d['envs'], d['l2_flushed_between_launches'], round(d['ms_step_kernel_median']*1e3,1), 'us', round(d['frac_algorithmic'],3))
    else: print('$1 $2', {k: round(v,4) if isinstance(v,float) else v for k,v in d.items() if k!='case'})"
done
