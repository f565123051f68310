#!/bin/bash
# round 2, job z: static first ticket, deeper taper (LVB_GPU_TAPER=2)
R=${GRAFT_REPO_ROOT:-/root/repo}; O=$R/gpurun_out; mkdir -p $O; cd $R
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_configs.py -m gpu -q -x > $O/r02z_pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $O/r02z_pytest_gpu.log
timeout 900 python tools/sweep_store.py --stream 256 --taper 1,2,1,2 --lag=-1 --warps 80,96,112 --shards 1 --reps 80 > $O/r02z_sweep.jsonl 2> $O/r02z_sweep.err
timeout 600 python tools/sweep_store.py --stream 128 --taper 1,2 --lag=-1 --warps 96 --shards 8 --reps 80 >> $O/r02z_sweep.jsonl 2>> $O/r02z_sweep.err
LVB_GPU_TAPER=2 timeout 300 python tools/trace_store.py 1 > $O/r02z_trace.log 2>&1
tail -n 3 $O/r02z_pytest_gpu.log; grep kernel_ms $O/r02z_sweep.jsonl | python -c "
import sys,json
for l in sys.stdin:
    d=json.loads(l); print(d['shards'], d['stream'], 'taper', d['taper'], 'tw', d['target_warps_per_sm'], d['kernel_ms'], d['length_ok'])"
grep "shards\|level\|flight" $O/r02z_trace.log | cut -c1-260
