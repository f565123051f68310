#!/bin/bash
# round 2, job j: SA probes — one chain per driver group (own thread, context, stream), device_open timing
R=${GRAFT_REPO_ROOT:-/root/repo}; O=$R/gpurun_out; mkdir -p $O; cd $R
nproc > $O/r02j_sa_probe.log
for spec in "8 8 4" "8 8 8" "16 16 8" "16 16 16" "64 16 4" "64 16 16" "64 32 32"; do set -- $spec
  echo "== chains $1 threads $2 groups $3" >> $O/r02j_sa_probe.log
  LVB_GPU_PROFILE=1 timeout 300 tools/sa_probe.sh 1000 200000 0.02 $1 3000 -a 1 -p $2 -G $3 2>&1 | grep "Wall\|All rep\|Host time\|device_open" >> $O/r02j_sa_probe.log
done
cat $O/r02j_sa_probe.log
