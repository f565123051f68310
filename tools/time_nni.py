"""dev tool: exhaustive NNI neighbourhood of one tree — lvb_gpu_nni_neighbourhood (up-pass + four-row
kernel) against the same 2(n-3) candidates scored as dirty paths by lvb_gpu_score_batch.
usage: python tools/time_nni.py [n nwords]   (default 2000 62500 = 2000 taxa x 1M sites)"""
import json, sys, time
sys.path.insert(0, '.'); sys.path.insert(0, 'tests')
import numpy as np
from lvb_b200 import Engine, Topology, synthetic
from test_gpu_parity import nni_neighbourhood

n, nwords = (int(sys.argv[1]), int(sys.argv[2])) if len(sys.argv) > 2 else (2000, 62500)
enc = synthetic.random_packed(n, nwords, 1)
l, r, root = synthetic.random_topology(n, 3)
e = Engine(0); e.upload(enc, max_trees=2); t, s = e.tree_create(), e.tree_create()
topo = Topology(l, r, root)
base = e.commit(t, topo)
got = e.nni_neighbourhood(t, s, topo)
reps = 5
e.synchronize(); t0 = time.perf_counter()
for _ in range(reps):
    got = e.nni_neighbourhood(t, s, topo)
dt_call = (time.perf_counter() - t0) / reps
kernel_ms = e.last_kernel_ms()
cands = nni_neighbourhood(l, r, root, n)
trees = [t] * len(cands)
want = e.score_batch(trees, cands)
assert np.array_equal(got, want), "neighbourhood != incremental path"
e.kernel_time(reset=True)
for _ in range(3):
    e.rerun_last_batch()
e.synchronize()
ms, k = e.kernel_time()
rows = 6 * (n - 3)  # A, B, C, D, low(u), up(u) per branch (operand rows; neighbouring branches share them through L2)
print(json.dumps({
    "shape": [n, nwords * 16], "candidates": len(cands), "current_length": int(base), "best_delta": int(got.min() - base),
    "nni_call_ms": dt_call * 1e3, "nni_call_candidates_per_s": len(cands) / dt_call,
    "four_row_kernel_ms": kernel_ms, "four_row_kernel_GBps": rows * nwords * 8 / kernel_ms / 1e6,
    "dirty_path_kernel_ms": ms / k, "dirty_path_candidates_per_s_device": len(cands) / (ms / k) * 1e3,
    "identical": True}))
e.close()
