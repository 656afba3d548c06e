#!/bin/bash
TAG=$1
mkdir -p gpurun_out
python profiles/scripts/mode_a_once.py 1048576 > gpurun_out/${TAG}_modeA_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:step_kernel -s 6 -c 1 -f -o gpurun_out/${TAG}_step python profiles/scripts/mode_a_once.py 1048576 > gpurun_out/${TAG}_ncu3.log 2>&1
tail -2 gpurun_out/${TAG}_ncu3.log
