#!/bin/bash
bash profiles/scripts/job_ab.sh "deep-rl-battlespace_b200/battlespace/libbattlespace.so $1"
python -m pytest tests/test_cuda_vs_oracle.py tests/test_full_size_oracle.py tests/test_rollout_mappings.py tests/test_cuda_golden.py -m gpu -x -q 2>&1 | tail -2
