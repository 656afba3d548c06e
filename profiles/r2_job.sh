#!/bin/bash
# One GPU-box job for round 2 (single GPU): A/B of library variants, the whole GPU test suite, smoke, the bench
# line + reference arm, the secondary configs, and the ncu passes of B200_PROFILING.md on the same short commands.
# usage: profiles/r2_job.sh TAG "ab/lib_a.so ab/lib_b.so ..."     (the in-tree library is the product build)
set -u
mkdir -p gpurun_out
TAG=$1; LIBS=${2:-}
for lib in $LIBS; do
  for rep in 1 2; do
    BATTLESPACE_LIB=$PWD/$lib python bench.py --steps 200 --no-cpu-baseline --no-e2e --no-sustained 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('$lib', round(d['value']/1e9,3), 'e9', round(d['ms_per_step'],4),'ms', d['clocks'])" | tee -a gpurun_out/${TAG}_ab.txt
  done
done
python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/${TAG}_pytest_gpu.log
tail -4 gpurun_out/${TAG}_pytest_gpu.log
python -c 'import __graft_entry__ as g; g.smoke()' > gpurun_out/${TAG}_smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/${TAG}_smoke.log
python bench.py > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err; echo "bench rc=$?" >> gpurun_out/${TAG}_bench.err
python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/${TAG}_bench_ref.json 2>> gpurun_out/${TAG}_bench.err
python profiles/bench_configs.py modeA modeAloops cfg2n cfg4 cfg5n > gpurun_out/${TAG}_configs.jsonl 2> gpurun_out/${TAG}_configs.err
CMD="python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-e2e --no-sustained"
$CMD > gpurun_out/${TAG}_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/${TAG}_launches.csv $CMD > gpurun_out/${TAG}_ncu1.log 2>&1
$CMD > gpurun_out/${TAG}_plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:rollout_kernel -s 3 -c 1 -f -o gpurun_out/${TAG}_rollout $CMD > gpurun_out/${TAG}_ncu2.log 2>&1
# Mode A at 1,048,576 envs: the step kernel from DRAM (working set 1.1 GB)
python profiles/scripts/mode_a_once.py 1048576 > gpurun_out/${TAG}_modeA_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:step_kernel -s 6 -c 1 -f -o gpurun_out/${TAG}_step python profiles/scripts/mode_a_once.py 1048576 > gpurun_out/${TAG}_ncu3.log 2>&1
# actor kernel (config 5 share)
python profiles/scripts/actor_once.py > gpurun_out/${TAG}_actor_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:actor_kernel -s 3 -c 1 -f -o gpurun_out/${TAG}_actor python profiles/scripts/actor_once.py > gpurun_out/${TAG}_ncu4.log 2>&1
tail -2 gpurun_out/${TAG}_smoke.log; cat gpurun_out/${TAG}_bench.json; cat gpurun_out/${TAG}_configs.jsonl | cut -c1-600
