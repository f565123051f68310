"""One rank of tests/test_gpu_configs.py::test_nccl_two_ranks_site_blocks: rank r owns its word
block of every row on GPU r; the NCCL unique id travels through a file (what MPI or
torch.distributed would do in a host program).  Prints one JSON line."""
import json
import os
import sys
import time

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, HERE)


def main():
    rank, world, td = int(sys.argv[1]), int(sys.argv[2]), sys.argv[3]
    import bench
    from lvb_b200 import Engine, Topology, sharding, synthetic

    n, m = 300, 60_000
    nwords = (m + 15) // 16
    enc = synthetic.packed(n, m, seed=5)
    left, right, root = synthetic.random_topology(n, seed=9)
    topo = Topology(left, right, root)
    cands, index = bench.nni_candidates(left, right, root, n, 16, seed=2)
    idf = os.path.join(td, "nccl_id")
    if rank == 0:
        uid = Engine.comm_unique_id()
        with open(idf + ".tmp", "wb") as f:
            f.write(uid)
        os.replace(idf + ".tmp", idf)
    else:
        t0 = time.time()
        while not os.path.exists(idf):
            if time.time() - t0 > 120:
                raise SystemExit("no NCCL id from rank 0")
            time.sleep(0.05)
        uid = open(idf, "rb").read()
    eng = Engine(rank)
    w0, wc = sharding.word_block(nwords, rank, world)
    eng.upload(enc, word_begin=w0, word_count=wc, max_trees=2)
    eng.comm_init(world, rank, uid)
    t, scratch = eng.tree_create(), eng.tree_create()
    out = {"rank": rank, "words": wc}
    out["commit"] = int(eng.commit(t, topo))
    out["changes"] = eng.changes(t)[n:].tolist()
    out["score"] = eng.score_batch([t] * len(cands), cands).tolist()
    nn = eng.nni_neighbourhood(t, scratch, topo)
    out["nni_at_index"] = nn[np.asarray(index)].tolist()
    eng.close()
    print(json.dumps(out))


if __name__ == "__main__":
    main()
