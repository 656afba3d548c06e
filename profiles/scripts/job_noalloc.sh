#!/bin/bash
# bs_rollout_actor_static: its test alone in a fresh process, then the whole GPU suite (it runs last), then the sanitizer job.  usage: job_noalloc.sh TAG
TAG=$1; mkdir -p gpurun_out
python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q > gpurun_out/${TAG}_noalloc_alone.log 2>&1; echo "alone rc=$?" | tee -a gpurun_out/${TAG}_noalloc_alone.log
tail -5 gpurun_out/${TAG}_noalloc_alone.log
bash profiles/scripts/job_tests.sh $TAG
bash profiles/scripts/job_sanitizer.sh $TAG
head -40 gpurun_out/${TAG}_sanitizer_memcheck.log
