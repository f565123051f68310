"""Test infrastructure: deterministic_hillclimb (Solve.c:64-142) from the same random start tree,
in a process of its own (the reference keeps file-static buffers sized by its first alignment),
printing one JSON line.
  --ref    the UNMODIFIED reference through oracle/_ref (the checker)
  --mock   the host layer with tests/mock_gpu.c answering lvb_gpu_* from the CPU oracle
  (none)   the host layer on the real library and a GPU
usage: run_hillclimb.py [--ref|--mock] <npz:key | synthetic:n:m:p> seed
"""
import ctypes as C
import hashlib
import json
import os
import sys
import tempfile

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, ROOT)
sys.path.insert(0, HERE)


def main():
    args = sys.argv[1:]
    mode = "gpu"
    if args[0] in ("--ref", "--mock"):
        mode = args[0][2:]
        args = args[1:]
    import numpy as np

    seed = int(args[1])
    if args[0].startswith("synthetic:"):  # synthetic:n:m:p — lvb_b200.synthetic alignment, generator seed 1
        from lvb_b200 import synthetic

        _, n_, m_, p_ = args[0].split(":")
        rows = [bytes(r) for r in synthetic.text_rows(synthetic.states(int(n_), int(m_), 1, float(p_)))]
    else:
        name, key = args[0].split(":")
        rows = [bytes(r) for r in np.load(os.path.join(HERE, "golden", name))[key]]
    import time

    t_start = time.perf_counter()
    with tempfile.TemporaryDirectory() as d:
        out = os.path.join(d, "outtree")
        if mode == "ref":
            import oracles

            msa = oracles.RefMSA.from_rows(rows)
            oracles.ref().ref_rinit(seed)
            t = msa.tree().random()
            res = (C.c_long * 3)()
            oracles.ref().ref_deterministic_hillclimb(msa.h, t.t, t.root, out.encode(), res)
            doc = dict(length=int(res[0]), iterations=int(res[1]), trees=int(res[2]))
        else:
            if mode == "mock":
                from run_host_search import build_mock

                C.CDLL(build_mock(), mode=C.RTLD_GLOBAL)
            from lvb_b200 import host

            msa = host.HostMSA.from_rows(rows)
            msa.rinit(seed)
            t = msa.tree().random()
            doc = t.hillclimb(0, out)
        doc["seconds"] = time.perf_counter() - t_start  # alignment set-up + start tree + hill climb + output
        doc["outtree_md5"] = hashlib.md5(open(out, "rb").read()).hexdigest()
    print(json.dumps(doc))


if __name__ == "__main__":
    main()
