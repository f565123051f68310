#!/bin/bash
# SA replicates of BASELINE.json configs[4] spread over the GPUs of one box by one lvb_gpu process
# (run under gpurun --gpus N): 64 chains of 1000 x 200k, -D 1/2/4/8 as far as the box goes.
R=${GRAFT_REPO_ROOT:-/root/repo}
ng=$(nvidia-smi -L | wc -l); echo "gpus $ng, host cores $(nproc)"
for d in 1 2 4 8; do
  [ $d -le $ng ] || continue
  echo "== config5 64 chains x ${1:-3000} rearrangements, -D $d"
  $R/tools/sa_probe.sh 1000 200000 0.02 64 ${1:-3000} -a 1 -D $d 2>&1 | grep -v Replicate | grep "Wall\|All rep\|Host time"
done
