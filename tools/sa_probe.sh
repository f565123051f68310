#!/bin/bash
# usage: tools/sa_probe.sh n m p R x [extra flags]  — synthetic alignment, SA for x rearrangements, prints the result block tail
set -e
n=$1; m=$2; p=$3; R=$4; x=$5; shift 5
d=$(mktemp -d)
python - <<PY
import sys; sys.path.insert(0,'$GRAFT_REPO_ROOT' if '$GRAFT_REPO_ROOT' else '/root/repo')
from lvb_b200 import synthetic
synthetic.write_phylip('$d/infile', synthetic.states($n,$m,1,$p))
PY
cd $d && ${GRAFT_REPO_ROOT:-/root/repo}/lvb_b200/bin/lvb_gpu -s 12345 -R $R -x $x "$@" | tail -12
