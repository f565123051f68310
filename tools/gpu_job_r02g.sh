#!/bin/bash
# round 2, job g: tapered level-0 bins (guided self-scheduling) A/B on one box
R=${GRAFT_REPO_ROOT:-/root/repo}; O=$R/gpurun_out; mkdir -p $O; cd $R
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_configs.py -m gpu -q -x > $O/r02g_pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $O/r02g_pytest_gpu.log
timeout 900 python tools/sweep_store.py --stream 128,256 --taper 0,1 --lag=-1 --warps 96,128,192,256,384 --shards 1,8 --reps 40 > $O/r02g_sweep.jsonl 2> $O/r02g_sweep.err
tail -n 5 $O/r02g_pytest_gpu.log; tail -n 3 $O/r02g_sweep.err
