#!/bin/bash
# A/B of library builds on the GPU box: parity subset on the LAST library given, then a short bench of each.
# usage: profiles/ab_libs.sh tag lib1.so lib2.so ...     (libraries are paths under ab/)
TAG=$1; shift
mkdir -p gpurun_out
LAST="${@: -1}"
BATTLESPACE_LIB=$PWD/$LAST timeout 900 python -m pytest tests/test_cuda_vs_oracle.py tests/test_cuda_golden.py -m gpu -x -q 2>&1 | tail -3 | tee gpurun_out/${TAG}_parity.log
for lib in "$@"; do
  for rep in 1 2; do
  BATTLESPACE_LIB=$PWD/$lib python bench.py --steps 200 --no-cpu-baseline --no-e2e 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('$lib', round(d['value']/1e9,3), 'e9', round(d['ms_per_step'],4),'ms', d['clocks'])" | tee -a gpurun_out/${TAG}_ab.txt
  done
done
