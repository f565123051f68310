#!/bin/bash
# round 2, job r: level release first in the bookkeeper; lag order again: parity, time line, sweep
R=${GRAFT_REPO_ROOT:-/root/repo}; O=$R/gpurun_out; mkdir -p $O; cd $R
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_configs.py -m gpu -q -x > $O/r02r_pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $O/r02r_pytest_gpu.log
timeout 300 python tools/trace_store.py 1 > $O/r02r_trace.log 2>&1
timeout 300 python tools/trace_store.py 8 >> $O/r02r_trace.log 2>&1
timeout 900 python tools/sweep_store.py --stream 128,256 --taper 1 --lag=-1,4,12,32 --warps 64,96,128,192 --shards 1,8 --reps 40 > $O/r02r_sweep.jsonl 2> $O/r02r_sweep.err
tail -n 4 $O/r02r_pytest_gpu.log; cat $O/r02r_trace.log; tail -3 $O/r02r_sweep.err
