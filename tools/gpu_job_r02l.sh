#!/bin/bash
# round 2, job l: SA probe — CTA width of the small-batch kernels (LVB_GPU_BLOCK) at 8 chains / 8 groups and 1 chain
R=${GRAFT_REPO_ROOT:-/root/repo}; O=$R/gpurun_out; mkdir -p $O; cd $R
: > $O/r02l_sa_probe.log
for blk in 256 128 64 32; do
for spec in "8 8 8" "1 1 1"; do set -- $spec
  echo "== block $blk chains $1 threads $2 groups $3" >> $O/r02l_sa_probe.log
  LVB_GPU_BLOCK=$blk LVB_GPU_PROFILE=1 timeout 300 tools/sa_probe.sh 1000 200000 0.02 $1 3000 -a 1 -p $2 -G $3 2>&1 | grep "Wall\|Host time" >> $O/r02l_sa_probe.log
done; done
echo "== config 3 shape, one chain" >> $O/r02l_sa_probe.log
for blk in 256 64; do LVB_GPU_BLOCK=$blk timeout 300 tools/sa_probe.sh 500 100000 0.02 1 20000 -a 1 -p 4 2>&1 | grep "Wall\|Host time" >> $O/r02l_sa_probe.log; done
cat $O/r02l_sa_probe.log
