#!/bin/bash
# SA probe on the GPU box, config 5 (64 chains of 1000 x 200k on one GPU): driver groups per GPU
R=${GRAFT_REPO_ROOT:-/root/repo}
f='^  [A-Z][a-z]* [a-z]*:\|^Search\|^$\|All topo\|Replicate\|0 calls'
run() { echo "== config5 $*"; $R/tools/sa_probe.sh 1000 200000 0.02 64 3000 -a 1 "$@" 2>&1 | grep -v "$f" | tail -3; }
run
run -G 1
run -G 2
run -G 4
run -G 8
