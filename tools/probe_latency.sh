#!/bin/bash
# single-chain round latency on the GPU box: batches inside the kernel parameters (default) against staged batches
R=${GRAFT_REPO_ROOT:-/root/repo}
f='^  [A-Z][a-z]* [a-z]*:\|^Search\|^$\|All topo\|Replicate'
run() { echo "== $*"; env "${ENVV[@]}" $R/tools/sa_probe.sh "$@" 2>&1 | grep -v "$f" | tail -3; }
for e in "A=1" "LVB_GPU_DEEP=0" "LVB_GPU_DEEP=0 LVB_GPU_INLINE=0 LVB_GPU_FUSE=0"; do
  ENVV=($e); echo "#### $e"
  run 100 984 0.05 1 60000 -a 1
  run 100 10000 0.05 1 40000 -a 1
  run 500 100000 0.02 1 20000 -a 1
done
