"""Test infrastructure: run the host layer's search (lvb_b200/host, GetSolnBatched) in a process of
its own and print one JSON line.  With --mock the lvb_gpu_* entry points are answered by
tests/mock_gpu.c (CPU oracle) — for the `-m "not gpu"` suite; without it the real
liblvbgpu.so and a GPU are used.

usage: run_host_search.py [--mock] <alignment.npz key | path.phy> seed algorithm [speculation] [replicates] [max_rearr] [host_threads] [groups]
"""
import ctypes as C
import hashlib
import json
import os
import subprocess
import sys
import tempfile

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, ROOT)
sys.path.insert(0, HERE)


def build_mock() -> str:
    out_dir = os.path.join(HERE, "_build")
    os.makedirs(out_dir, exist_ok=True)
    so = os.path.join(out_dir, "libmockgpu.so")
    srcs = [os.path.join(HERE, "mock_gpu.c"), os.path.join(ROOT, "oracle", "fitch_oracle.c")]
    if not os.path.exists(so) or any(os.path.getmtime(s) > os.path.getmtime(so) for s in srcs):
        tmp = so + f".{os.getpid()}.tmp"
        subprocess.run(["gcc", "-O2", "-std=c11", "-shared", "-fPIC", "-o", tmp] + srcs, check=True)
        os.replace(tmp, so)
    return so


def md5(path):
    return hashlib.md5(open(path, "rb").read()).hexdigest()


def main():
    args = sys.argv[1:]
    mock = False
    if args and args[0] == "--mock":
        mock = True
        args = args[1:]
        C.CDLL(build_mock(), mode=C.RTLD_GLOBAL)
    import numpy as np
    from lvb_b200 import host

    src, seed, algo = args[0], int(args[1]), int(args[2])
    spec = int(args[3]) if len(args) > 3 else 0
    reps = int(args[4]) if len(args) > 4 else 1
    max_rearr = int(args[5]) if len(args) > 5 else 0
    threads = int(args[6]) if len(args) > 6 else 1
    groups = int(args[7]) if len(args) > 7 else 0
    if src.endswith(".phy"):
        msa = host.HostMSA.load(src)
    else:
        name, key = src.split(":")
        g = np.load(os.path.join(HERE, "golden", name))
        msa = host.HostMSA.from_rows([bytes(r) for r in g[key]])
    with tempfile.TemporaryDirectory() as d:
        tsv, out = os.path.join(d, "t.tsv"), os.path.join(d, "outtree")
        res = msa.search(seed, algorithm=algo, speculation=spec, replicates=reps, max_rearrangements=max_rearr, host_threads=threads,
                         groups=groups, tsv_path=tsv, outtree_path=out)
        line = {"mock": mock, "results": res, "tsv_md5": md5(tsv), "outtree_md5": md5(out),
                "tsv_head": open(tsv).read().split("\n")[:41],
                "outtrees": [md5(out + (f".r{r}" if r else "")) for r in range(reps)]}
    print(json.dumps(line))


if __name__ == "__main__":
    main()
