#!/bin/bash
# e2e A/B: python -m pytest host rollout test, then bench e2e for each env setting
true
for setting in "$@"; do
  env $setting python bench.py --steps 50 --no-cpu-baseline | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('$setting', round(d['value']/1e9,3), 'e2e', round(d['e2e']['value']/1e9,3))"
done
