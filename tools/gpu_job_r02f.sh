#!/bin/bash
# round 2, job f: streaming store kernel with warp-0 bookkeeping: parity, sweep, per-step latency probe (one piece, 4 CTAs)
R=${GRAFT_REPO_ROOT:-/root/repo}; O=$R/gpurun_out; mkdir -p $O; cd $R
timeout 1200 python -m pytest tests/test_gpu_parity.py tests/test_gpu_configs.py -m gpu -q -x > $O/r02f_pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $O/r02f_pytest_gpu.log
timeout 900 python tools/sweep_store.py --stream 128,256 --lag=-1 --warps 96,128,192,256,384,768 --shards 1,8 --reps 40 > $O/r02f_sweep.jsonl 2> $O/r02f_sweep.err
timeout 300 python tools/sweep_store.py --stream 128,256,0 --pf 7 --block 256 --lag=-1 --warps 1 --shards 64 --reps 20 > $O/r02f_latency.jsonl 2>> $O/r02f_sweep.err
tail -n 5 $O/r02f_pytest_gpu.log; tail -n 3 $O/r02f_sweep.err; cat $O/r02f_latency.jsonl
