#!/bin/bash
# compute-sanitizer memcheck + racecheck over every kernel at small ragged sizes.  usage: job_sanitizer.sh TAG
TAG=$1; mkdir -p gpurun_out
python profiles/scripts/exercise_all_kernels.py > gpurun_out/${TAG}_exercise_plain.log 2>&1; tail -1 gpurun_out/${TAG}_exercise_plain.log
for tool in memcheck racecheck; do
  timeout 600 compute-sanitizer --tool $tool --error-exitcode 9 python profiles/scripts/exercise_all_kernels.py > gpurun_out/${TAG}_sanitizer_$tool.log 2>&1
  echo "$tool rc=$?" | tee -a gpurun_out/${TAG}_sanitizer_$tool.log
  grep -E "ERROR SUMMARY|RACECHECK SUMMARY|Error:|unable|not supported|denied" gpurun_out/${TAG}_sanitizer_$tool.log | head -5
done
