#!/bin/bash
# SA host-side probe on the GPU box: per-stage host profile (LVB_GPU_PROFILE) under several
# thread / group settings.  usage: tools/probe_host.sh > gpurun_out/probe_host.log
R=${GRAFT_REPO_ROOT:-/root/repo}
export LVB_GPU_PROFILE=${LVB_GPU_PROFILE:-1}
f='^  [A-Z][a-z]* [a-z]*:\|^Search\|^$\|All topo\|Replicate'
run() { echo "== config5 $*"; env "${ENVV[@]}" $R/tools/sa_probe.sh 1000 200000 0.02 64 3000 -a 1 "$@" 2>&1 | grep -v "$f" | tail -4; }
ENVV=(A=1); run
ENVV=(A=1); run -p 8
ENVV=(A=1); run -K 2
ENVV=(A=1); run -G 2
ENVV=(LVB_GPU_COMPILE_PAR=4); run
