#!/bin/bash
TAG=$1; mkdir -p gpurun_out
python bench.py > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err; echo "bench rc=$?"
python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/${TAG}_bench_ref.json 2>> gpurun_out/${TAG}_bench.err
python profiles/bench_configs.py cfg4 cfg5n > gpurun_out/${TAG}_configs.jsonl 2>/dev/null
CMD="python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-e2e --no-sustained"
$CMD > gpurun_out/${TAG}_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/${TAG}_launches.csv $CMD > gpurun_out/${TAG}_ncu1.log 2>&1
$CMD > gpurun_out/${TAG}_plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:rollout_kernel -s 3 -c 1 -f -o gpurun_out/${TAG}_rollout $CMD > gpurun_out/${TAG}_ncu2.log 2>&1
python -c 'import __graft_entry__ as g; g.smoke()' 2>&1 | tail -1
cut -c1-300 gpurun_out/${TAG}_bench.json; cut -c1-260 gpurun_out/${TAG}_configs.jsonl
