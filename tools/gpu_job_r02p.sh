#!/bin/bash
# round 2, job p: in-kernel time line of a whole-tree store launch (full width and 1/8), SA probe with groups bounded by threads, ingest
R=${GRAFT_REPO_ROOT:-/root/repo}; O=$R/gpurun_out; mkdir -p $O; cd $R
timeout 300 python tools/trace_store.py 1 > $O/r02p_trace.log 2>&1
timeout 300 python tools/trace_store.py 8 >> $O/r02p_trace.log 2>&1
: > $O/r02p_sa_probe.log
for spec in "8 4" "8 8" "64 16"; do set -- $spec
  echo "== chains $1 threads $2 (groups automatic)" >> $O/r02p_sa_probe.log
  LVB_GPU_PROFILE=1 timeout 300 taskset -c 0-$(( $2 - 1 )) tools/sa_probe.sh 1000 200000 0.02 $1 3000 -a 1 -p $2 2>&1 | grep "Wall\|All rep\|Host time\|ERROR" >> $O/r02p_sa_probe.log
done
timeout 600 python tools/time_ingest.py > $O/r02p_ingest.json 2> $O/r02p_ingest.err
cat $O/r02p_trace.log $O/r02p_sa_probe.log $O/r02p_ingest.json; tail -3 $O/r02p_ingest.err
