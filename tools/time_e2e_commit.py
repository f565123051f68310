"""dev tool: where the end-to-end time of one full-pass lvb_gpu_tree_commit goes (2000 x 1M by default):
python + ctypes, host compile, layout + H2D, launch, wait.  LVB_GPU_PROFILE=1 prints the C-side split."""
import os, sys, time
os.environ.setdefault('LVB_GPU_PROFILE', '1')
sys.path.insert(0, '.')
from lvb_b200 import Engine, Topology, synthetic

n, nwords = (int(sys.argv[1]), int(sys.argv[2])) if len(sys.argv) > 2 else (2000, 62500)
steps = 200
REPS = int(os.environ.get('REPS', '1'))  # passes over the 200 topologies (more = warm-up amortised in the C-side profile)
enc = synthetic.random_packed(n, nwords, 1)
e = Engine(0); e.upload(enc, max_trees=2); t = e.tree_create()
topos = [Topology(*synthetic.random_topology(n, seed=20 + s)) for s in range(steps)]
for tp in topos[:5]:
    e.commit(t, tp)
e.kernel_time(reset=True)
t0 = time.perf_counter()
for rep in range(REPS):
    for tp in topos:
        e.commit(t, tp)
wall = (time.perf_counter() - t0) / REPS
ms, k = e.kernel_time()
print('e2e %.1f us per commit; main kernel %.1f us per launch (%d launches)' % (wall / steps * 1e6, ms / max(1, k) * 1e3, k))
t0 = time.perf_counter()
for i in range(steps):
    e.rerun_last_batch()
e.synchronize()
print('rerun (device-resident) %.1f us per pass' % ((time.perf_counter() - t0) / steps * 1e6))
t0 = time.perf_counter()
for tp in topos:
    tp.as_struct()
print('python as_struct %.1f us' % ((time.perf_counter() - t0) / steps * 1e6))
e.close()

# host compile alone, steady state (the same call lvb_gpu_tree_commit makes inside)
import ctypes as C
import numpy as np
from lvb_b200.gpu import load_library
lib = load_library()
out = np.zeros((4 * n, 4), dtype=np.int32); pcs = np.zeros((4 * n, 3), dtype=np.int32)
k, sl, npc = C.c_int64(0), C.c_int64(0), C.c_int64(0)
structs = [tp.as_struct() for tp in topos]
for what, seq in (('a new topology every call', structs), ('the same topology every call', [structs[0]] * steps)):
    t0 = time.perf_counter()
    for ts in seq:
        lib.lvb_gpu_compile_schedule(n, C.byref(ts), 1, out.ctypes.data, 4 * n, C.byref(k), C.byref(sl))
    print('lvb_gpu_compile_schedule (persistent compiler), %s: %.1f us per tree' % (what, (time.perf_counter() - t0) / steps * 1e6))
