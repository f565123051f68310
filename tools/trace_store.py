#!/usr/bin/env python
"""dev tool: where the time of ONE whole-tree store launch goes (LVB_GPU_TRACE=1: six device time stamps per work
item of the streaming kernel).  usage: trace_store.py [shards]  (word block of 1/shards of 2000 x 1M)"""
import ctypes as C, os, sys
os.environ["LVB_GPU_TRACE"] = "1"
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from lvb_b200 import Engine, Topology, synthetic, compile_pieces, load_library

shards = int(sys.argv[1]) if len(sys.argv) > 1 else 1
n, nwords = 2000, 62500
enc = synthetic.random_packed(n, nwords, seed=1)
left, right, root = synthetic.random_topology(n, seed=3)
eng = Engine(0)
wc = nwords // shards
eng.upload(enc, word_begin=0, word_count=wc, max_trees=1)
t = eng.tree_create()
eng.commit(t, Topology(left, right, root))
for _ in range(3):
    eng.rerun_last_batch()
eng.synchronize()
lib = load_library()
lib.lvb_gpu_trace_download.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.POINTER(C.c_int64), C.POINTER(C.c_int64)]
ni, ncb = C.c_int64(0), C.c_int64(0)
assert lib.lvb_gpu_trace_download(eng.ctx, None, 0, C.byref(ni), C.byref(ncb)) == 0
tr = np.zeros((ni.value, 6), dtype=np.uint64)
assert lib.lvb_gpu_trace_download(eng.ctx, tr.ctypes.data, ni.value, C.byref(ni), C.byref(ncb)) == 0
n_cb = ncb.value
cols = (wc + 1) // 2
stride = (cols + 7) // 8 * 8
want = max(1, 148 * 96 // ((stride + 31) // 32))
piece_size = max(64, -(-1997 // want))
instr, pieces, slots = compile_pieces(n, Topology(left, right, root), piece_size)
levels = np.repeat(pieces[:, 2], n_cb)
sizes = np.repeat(pieces[:, 1], n_cb)
assert len(levels) == ni.value, (len(levels), ni.value)
t0 = tr[:, 0].min()
us = (tr.astype(np.int64) - int(t0)) / 1e3
print(f"shards {shards}: {len(pieces)} pieces x {n_cb} column blocks = {ni.value} items, piece_size {piece_size}, "
      f"last stamp {us.max():.1f} us after the first ticket")
names = ["ticket", "decoded", "published", "first operand", "last step", "books kept"]
for k in range(6):
    print(f"  {names[k]:>14}: first {us[:, k].min():7.1f}  median {np.median(us[:, k]):7.1f}  last {us[:, k].max():7.1f}")
for L in sorted(set(levels.tolist())):
    m = levels == L
    dur = us[m, 4] - us[m, 3]
    wait = us[m, 2] - us[m, 1]
    print(f"  level {L}: {m.sum():5d} items of {sizes[m].min()}..{sizes[m].max()} instr; published {us[m, 2].min():7.1f}..{us[m, 2].max():7.1f}, "
          f"last step {us[m, 4].min():7.1f}..{us[m, 4].max():7.1f}; first operand -> last step: median {np.median(dur):6.1f} us "
          f"({np.median(dur / sizes[m]) * 1e3:5.0f} ns/instr), wait for the level below: median {np.median(wait):5.1f} max {wait.max():5.1f}; "
          f"ticket -> first operand median {np.median(us[m, 3] - us[m, 0]):5.1f}")
# how busy the device is over time: items between first operand and last step, in 10 us bins
edges = np.arange(0, us.max() + 10, 10)
busy = [(int(e), int(((us[:, 3] <= e) & (us[:, 4] > e)).sum())) for e in edges]
print("  items in flight at t (us):", " ".join(f"{e}:{b}" for e, b in busy))
eng.close()
