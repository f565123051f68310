#!/bin/bash
# round 2, job d: lagged work order of the streaming store kernel: parity (full-size + config shapes + search), sweep of
# lag / piece size / consumer threads, SA probes with more driver groups per GPU (smem-limit race fixed)
R=${GRAFT_REPO_ROOT:-/root/repo}; O=$R/gpurun_out; mkdir -p $O; cd $R
timeout 1200 python -m pytest tests/test_gpu_parity.py tests/test_gpu_configs.py -m gpu -q -x > $O/r02d_pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $O/r02d_pytest_gpu.log
timeout 900 python tools/sweep_store.py --stream 128,256 --lag=-1,0,16,48,1000 --warps 96,192,384,768 --shards 1,8 --reps 40 > $O/r02d_sweep.jsonl 2> $O/r02d_sweep.err
for spec in "8 8 1" "8 8 2" "8 8 4" "64 16 4" "64 16 8"; do set -- $spec
  echo "== chains $1 threads $2 groups $3" >> $O/r02d_sa_probe.log
  LVB_GPU_PROFILE=2 timeout 300 tools/sa_probe.sh 1000 200000 0.02 $1 3000 -a 1 -p $2 -G $3 2>&1 | grep -v "Replicate" >> $O/r02d_sa_probe.log
done
tail -n 5 $O/r02d_pytest_gpu.log; tail -n 3 $O/r02d_sweep.err; grep "Wall\|==" $O/r02d_sa_probe.log
