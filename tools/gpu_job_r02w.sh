#!/bin/bash
# round 2, job w: compute-sanitizer (memcheck, synccheck) over the store routes, score-and-keep and a short search; new tests;
# C-ABI call timing with and without torch in the process
R=${GRAFT_REPO_ROOT:-/root/repo}; O=$R/gpurun_out; mkdir -p $O; cd $R
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -q -x -k "keep or trace or routes or errors" > $O/r02w_pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $O/r02w_pytest_gpu.log
cat > /tmp/san.py <<'PY'
import sys, numpy as np
sys.path.insert(0, '.'); sys.path.insert(0, 'tests')
import oracles
from lvb_b200 import Engine, Topology, synthetic
sys.path.insert(0, 'tests')
from test_gpu_parity import _nni_of
n, m = 150, 9000
enc = synthetic.pack_states(synthetic.states(n, m, seed=5, p=0.1))
eng = Engine(0); eng.upload(enc, max_trees=3)
t = eng.tree_create(); s = eng.tree_create()
left, right, root = synthetic.random_topology(n, seed=30)
want, rows, chg = oracles.oracle_getplen(enc, left, right, root)
assert eng.commit(t, Topology(left, right, root)) == want          # dependent pieces: streaming kernel
for u in (n + 3, n + 40, n + 99):
    l2, r2, dirty = _nni_of(left, right, root, n, u, True)
    cand = Topology(l2, r2, root, dirty)
    lens, ticket = eng.score_keep([t, t], [cand, cand], [1, 0])     # peek finalize, spare bank
    assert lens[0] == lens[1] == eng.score_batch([t], [cand])[0]
    eng.accept_kept(ticket, [0])
    left, right = l2, r2
assert eng.score_batch([t], [Topology(left, right, root, [])])[0] == oracles.oracle_getplen(enc, left, right, root)[0]
nn = eng.nni_neighbourhood(t, s, Topology(left, right, root))
assert eng.commit(t, Topology(left, right, root)) == oracles.oracle_getplen(enc, left, right, root)[0]
eng.close(); print("sanitizer workload ok")
PY
timeout 900 compute-sanitizer --tool memcheck --error-exitcode 9 python /tmp/san.py > $O/r02w_memcheck.log 2>&1; echo "memcheck rc=$?" >> $O/r02w_memcheck.log
timeout 900 compute-sanitizer --tool synccheck --error-exitcode 9 python /tmp/san.py > $O/r02w_synccheck.log 2>&1; echo "synccheck rc=$?" >> $O/r02w_synccheck.log
timeout 300 python tools/profile_score.py > $O/r02w_score.log 2>&1
WITH_TORCH=1 timeout 300 python tools/profile_score.py >> $O/r02w_score.log 2>&1
tail -n 3 $O/r02w_pytest_gpu.log; tail -n 6 $O/r02w_memcheck.log; tail -n 6 $O/r02w_synccheck.log; grep "C-ABI call" $O/r02w_score.log
