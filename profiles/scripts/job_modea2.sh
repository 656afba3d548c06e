#!/bin/bash
for lib in deep-rl-battlespace_b200/battlespace/libbattlespace.so ab/lib_nopf.so ab/lib_pf2.so ab/lib_pf05.so; do
  BATTLESPACE_LIB=$PWD/$lib python profiles/bench_configs.py modeA 2>/dev/null | python -c "
import sys, json
for l in sys.stdin:
    d = json.loads(l)
    if 'envs' in d: print('$lib', d['envs'], d['l2_flushed_between_launches'], round(d['ms_step_kernel_median']*1e3,1), 'us', round(d['frac_algorithmic'],3))"
done
python -m pytest tests/test_cuda_golden.py tests/test_edges.py -m gpu -x -q 2>&1 | tail -1
