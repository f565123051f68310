"""dev tool: where the end-to-end time of one full-pass lvb_gpu_tree_commit goes (2000 x 1M by default):
python + ctypes, host compile, layout + H2D, launch, wait.  LVB_GPU_PROFILE=1 prints the C-side split."""
import os, sys, time
os.environ.setdefault('LVB_GPU_PROFILE', '1')
sys.path.insert(0, '.')
from lvb_b200 import Engine, Topology, synthetic

n, nwords = (int(sys.argv[1]), int(sys.argv[2])) if len(sys.argv) > 2 else (2000, 62500)
steps = 200
enc = synthetic.random_packed(n, nwords, 1)
e = Engine(0); e.upload(enc, max_trees=2); t = e.tree_create()
topos = [Topology(*synthetic.random_topology(n, seed=20 + s)) for s in range(steps)]
for tp in topos[:5]:
    e.commit(t, tp)
e.kernel_time(reset=True)
t0 = time.perf_counter()
for tp in topos:
    e.commit(t, tp)
wall = time.perf_counter() - t0
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
