#!/bin/bash
TAG=$1
mkdir -p gpurun_out
python -m pytest tests/test_actor_and_sample.py -m gpu -x -q 2>&1 | grep -E "assert|Error|error|>  |^E " | head -30
python profiles/scripts/actor_once.py > gpurun_out/${TAG}_actor_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:actor_kernel -s 3 -c 1 -f -o gpurun_out/${TAG}_actor python profiles/scripts/actor_once.py > gpurun_out/${TAG}_ncu4.log 2>&1
tail -2 gpurun_out/${TAG}_ncu4.log
