#!/bin/bash
# round 2, job ac (2 GPUs): the 2-rank site-block tests (peer-memory and NCCL routes) on the final code (row banks, score-and-keep)
R=${GRAFT_REPO_ROOT:-/root/repo}; O=$R/gpurun_out; mkdir -p $O; cd $R
timeout 900 python -m pytest tests/test_gpu_configs.py tests/test_zz_no_alloc.py -m gpu -q > $O/r02ac_pytest_gpu_2gpus.log 2>&1; echo "pytest rc=$?" >> $O/r02ac_pytest_gpu_2gpus.log
tail -n 6 $O/r02ac_pytest_gpu_2gpus.log
