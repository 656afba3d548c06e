#!/bin/bash
# A/B of library variants on the bench command (device-timed leg only)
# usage: job_ab.sh "ab/lib_a.so ab/lib_b.so ..." [extra env assignments]
for lib in $1; do
  for rep in 1 2; do
    env $2 BATTLESPACE_LIB=$PWD/$lib python bench.py --steps 100 --no-cpu-baseline --no-e2e --no-sustained 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('$lib', round(d['value']/1e9,3), 'e9', round(d['ms_per_step'],4),'ms', d['stats']['games'])"
  done
done
