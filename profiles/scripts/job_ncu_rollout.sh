#!/bin/bash
# ncu --set full of the fused rollout kernel of the bench command (one launch), after a plain run of the same command
# usage: job_ncu_rollout.sh TAG [kernel regex]
TAG=$1; K=${2:-rollout}
mkdir -p gpurun_out
CMD="python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-e2e --no-sustained"
$CMD > gpurun_out/${TAG}_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:$K -s 3 -c 1 -f -o gpurun_out/${TAG}_rollout $CMD > gpurun_out/${TAG}_ncu.log 2>&1
tail -3 gpurun_out/${TAG}_ncu.log
