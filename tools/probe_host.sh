#!/bin/bash
# SA host-side probe on the GPU box: per-stage host profile (LVB_GPU_PROFILE) under several
# thread / group settings.  usage: tools/probe_host.sh > gpurun_out/probe_host.log
R=${GRAFT_REPO_ROOT:-/root/repo}
export LVB_GPU_PROFILE=2
f='^  [A-Z][a-z]* [a-z]*:\|^Search\|^$\|All topo\|Replicate'
echo "== config5 -G 1 -p 1"; $R/tools/sa_probe.sh 1000 200000 0.02 64 3000 -a 1 -G 1 -p 1 2>&1 | grep -v "$f" | tail -12
echo "== config5 -G 1 -p 16"; $R/tools/sa_probe.sh 1000 200000 0.02 64 3000 -a 1 -G 1 -p 16 2>&1 | grep -v "$f" | tail -12
echo "== config3"; $R/tools/sa_probe.sh 500 100000 0.02 1 20000 -a 1 2>&1 | grep -v "$f" | tail -12
echo "== config2-ish 100x10000 40000 rearr"; $R/tools/sa_probe.sh 100 10000 0.05 1 40000 -a 1 2>&1 | grep -v "$f" |  tail -12
