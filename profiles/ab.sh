#!/bin/bash
# quick A/B on the GPU box: python -m pytest subset + short bench for each env-var setting given as args
python -m pytest tests/test_cuda_vs_oracle.py -m gpu -x -q -k "rollout or shard" 2>&1 | tail -2
for setting in "$@"; do
  env $setting python bench.py --steps 100 --no-cpu-baseline --no-e2e | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('$setting', round(d['value']/1e9,3), 'e9', round(d['ms_per_step'],4),'ms')"
done
