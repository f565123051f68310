#!/bin/bash
# Multi-GPU measurements on one box (run under gpurun --gpus 8): site blocks + NCCL all-reduce at
# 4 and 8 GPUs, replicas at 8, SA replicates spread over 1/2/4/8 GPUs by one lvb_gpu process.
R=${GRAFT_REPO_ROOT:-/root/repo}
O=$R/gpurun_out
mkdir -p $O
nproc > $O/scale_nproc.txt; nvidia-smi -L >> $O/scale_nproc.txt
tr() { python -m torch.distributed.run --nnodes=1 --nproc-per-node $1 --master-addr 127.0.0.1 --master-port $2 bench.py --gpus $1 "${@:3}"; }
cd $R
tr 8 29511 --steps 100 --warmup 5 --mode siteblock --no-cpu-baseline > $O/scale_siteblock_n8.json 2> $O/scale_siteblock_n8.err
CUDA_VISIBLE_DEVICES=0,1,2,3 tr 4 29512 --steps 100 --warmup 5 --mode siteblock --no-cpu-baseline > $O/scale_siteblock_n4.json 2> $O/scale_siteblock_n4.err
tr 8 29513 --steps 100 --warmup 5 --no-cpu-baseline > $O/scale_replicas_n8.json 2> $O/scale_replicas_n8.err
for d in 1 2 4 8; do
  echo "== config5 64 chains, -D $d"; tools/sa_probe.sh 1000 200000 0.02 64 3000 -a 1 -D $d 2>&1 | grep -v Replicate | tail -6
done > $O/scale_sa_config5.log 2>&1
tail -n 3 $O/scale_*.json $O/scale_*.err; cat $O/scale_sa_config5.log
