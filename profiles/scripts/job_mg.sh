#!/bin/bash
# Multi-GPU job (gpurun --gpus N): the NCCL test, BASELINE configs 4 and 5 in their multi-GPU shape, the concurrent
# pinned-H2D ceiling and the bench line, all under torch.distributed.run with one rank per GPU.
# usage: job_mg.sh TAG N
TAG=$1; N=$2
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29555"
python -m pytest tests/test_multi_gpu.py -m gpu -x -q > gpurun_out/${TAG}_pytest_mg_n$N.log 2>&1; tail -2 gpurun_out/${TAG}_pytest_mg_n$N.log
$TR profiles/bench_configs.py cfg4mg cfg5mg h2d > gpurun_out/${TAG}_cfg_n$N.jsonl 2> gpurun_out/${TAG}_cfg_n$N.err; echo "cfg rc=$?"
$TR bench.py --gpus $N --no-cpu-baseline > gpurun_out/${TAG}_bench_n$N.json 2> gpurun_out/${TAG}_bench_n$N.err; echo "bench rc=$?"
cut -c1-900 gpurun_out/${TAG}_cfg_n$N.jsonl; cut -c1-400 gpurun_out/${TAG}_bench_n$N.json
python - <<PY
import json
d=json.loads(open("gpurun_out/${TAG}_bench_n$N.json").read().strip().splitlines()[-1])
print("value", d["value"], "e2e", d["e2e"])
PY
