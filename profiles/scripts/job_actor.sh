#!/bin/bash
TAG=$1
mkdir -p gpurun_out
python -m pytest tests/test_actor_and_sample.py -m gpu -x -q 2>&1 | grep -E "^E  |passed|failed" | head -12
python profiles/bench_configs.py cfg5n cfg2n > gpurun_out/${TAG}_actor_cfg.jsonl 2>/dev/null; cut -c1-520 gpurun_out/${TAG}_actor_cfg.jsonl
