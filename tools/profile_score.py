#!/usr/bin/env python
"""dev tool for ncu: the score-path kernels at the bench configuration — 512 NNI candidates of a 2000 x 1M tree
(fitch_score_kernel), a small batch of 4 (fitch_pipe_kernel<false>), one accepted move (fitch_pipe_kernel<true>) and the
exhaustive NNI neighbourhood (up-pass pieces + nni_neighbourhood_kernel)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from lvb_b200 import Engine, Topology, synthetic
import bench

n, m = 2000, 1_000_000
enc = synthetic.random_packed(n, (m + 15) // 16, seed=1)
left, right, root = synthetic.random_topology(n, seed=3)
eng = Engine(0)
eng.upload(enc, max_trees=2)
t = eng.tree_create()
s = eng.tree_create()
print("length", eng.commit(t, Topology(left, right, root)))
cands, idx = bench.nni_candidates(left, right, root, n, 512, seed=5)
for _ in range(2):
    lens = eng.score_batch([t] * 512, cands)
print("score 512:", int(lens[0]), eng.last_kernel_ms())
import time
prep = eng.prepare_batch([t] * 512, cands)
eng.score_prepared(prep)
for label in ("no torch", "torch imported"):
    t0 = time.perf_counter()
    for _ in range(10):
        eng.score_prepared(prep)
    print(f"C-ABI call, 512 candidates, {label}: {(time.perf_counter() - t0) / 10 * 1e3:.3f} ms per call (kernel {eng.last_kernel_ms():.3f} ms)")
    if label == "no torch" and os.environ.get("WITH_TORCH"):
        import torch  # noqa: F401
    elif label == "no torch":
        break
print("score 4:", eng.score_batch([t] * 4, cands[:4]).tolist(), eng.last_kernel_ms())
print("commit of one candidate:", eng.commit(t, cands[0]), eng.last_kernel_ms())
nn = eng.nni_neighbourhood(t, s, Topology(cands[0].full[0], cands[0].full[1], root))
print("nni:", int(nn[0]), eng.last_kernel_ms())
eng.close()
