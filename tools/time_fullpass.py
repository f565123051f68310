"""dev tool: time the full-pass commit (default 2000 x 1M; or `n nwords`, e.g. 2000 7813 = one of 8 site blocks) for several settings"""
import os, sys, time
sys.path.insert(0, '.'); sys.path.insert(0, 'tests')
import numpy as np
from lvb_b200 import Engine, Topology, synthetic

n, nwords = (int(sys.argv[1]), int(sys.argv[2])) if len(sys.argv) > 2 else (2000, 62500)
enc = synthetic.random_packed(n, nwords, 1)
l, r, root = synthetic.random_topology(n, 3)
want = None
tws = [int(x) for x in os.environ.get('TWS', '128,192,256,384,512,768,1024').split(',')]
for ll in [int(x) for x in os.environ.get('LLS', '0,1').split(',')]:
    for tw in tws:
        for blk in [int(x) for x in os.environ.get('BLKS', '256').split(',')]:
            os.environ['LVB_GPU_LEVEL_LAUNCHES'] = str(ll)
            os.environ['LVB_GPU_TARGET_WARPS'] = str(148 * tw)
            os.environ['LVB_GPU_BLOCK'] = str(blk)
            e = Engine(0); e.upload(enc, max_trees=1); t = e.tree_create()
            L = e.commit(t, Topology(l, r, root))
            want = want or L
            assert L == want, (L, want)
            l0 = e.launches
            e.rerun_last_batch(); e.synchronize()
            per_call = e.launches - l0
            e.kernel_time(reset=True)
            for i in range(10): e.rerun_last_batch()
            e.synchronize()
            ms, k = e.kernel_time()
            per = ms / k
            L2 = e.commit(t, Topology(l, r, root))
            assert L2 == want
            print('level_launches %d target_warps 148*%4d block %3d launches/pass %d: %.3f ms/pass, algorithmic %.0f GB/s, %.3e updates/s' % (
                ll, tw, blk, per_call, per, (3 * n - 6) * 8 * nwords / per / 1e6, (n - 1) * nwords * 16 / per * 1e3), flush=True)
            e.close()
