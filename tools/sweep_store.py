#!/usr/bin/env python
"""Sweep the store kernel's knobs on one GPU: look-ahead (LVB_GPU_STORE_PF: 0 = register-pipelined
fitch_exec_kernel<true>, 2/3/5/7 = cp.async ring), piece size (LVB_GPU_TARGET_WARPS), CTA size
(LVB_GPU_BLOCK), on the full pass of one 2000 x 1M tree and on the word blocks a rank holds at
2/4/8 GPUs (site-block mode on ONE device: same kernel, same grid as on a rank).

Prints one JSON line per setting: kernel ms per pass (CUDA events around every launch, mean of
`reps` re-runs), stream ms per pass (events around the whole queue), the length (must not change).
"""
import argparse
import itertools
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--taxa", type=int, default=2000)
    ap.add_argument("--sites", type=int, default=1_000_000)
    ap.add_argument("--reps", type=int, default=30)
    ap.add_argument("--pf", default="0,2,3,5,7")
    ap.add_argument("--warps", default="96,192,384,768,1536")
    ap.add_argument("--block", default="256,128")
    ap.add_argument("--shards", default="1,8")
    ap.add_argument("--taper", default="1", help="LVB_GPU_TAPER values")
    ap.add_argument("--piece-min", default="64", help="LVB_GPU_PIECE_MIN values")
    ap.add_argument("--lag", default="-1", help="LVB_GPU_STREAM_LAG values (-1 = automatic)")
    ap.add_argument("--stream", default="0", help="LVB_GPU_STORE_STREAM values (consumer threads of the TMA streaming kernel; 0 = the other kernels)")
    args = ap.parse_args()

    import torch
    from lvb_b200 import Engine, Topology, synthetic

    n, m = args.taxa, args.sites
    nwords = (m + 15) // 16
    t0 = time.time()
    enc = synthetic.random_packed(n, nwords, seed=1)
    left, right, root = synthetic.random_topology(n, seed=3)
    topo = Topology(left, right, root)
    print(json.dumps({"setup_s": round(time.time() - t0, 1)}), flush=True)
    want = {}
    for shards in [int(x) for x in args.shards.split(",")]:
        wc = nwords // shards
        grid = []
        for st in args.stream.split(","):
            if int(st) > 0:
                grid += [(st, lg, tw, "256") for lg in args.lag.split(",") for tw in args.warps.split(",")]
            else:
                grid += [(st, pf, tw, blk) for pf, tw, blk in itertools.product(args.pf.split(","), args.warps.split(","), args.block.split(","))]
        grid = [(st, pf, tw, blk, tp, pm) for pm in args.piece_min.split(",") for tp in args.taper.split(",") for st, pf, tw, blk in grid]
        for st, pf, tw, blk, tp, pm in grid:
            os.environ["LVB_GPU_TAPER"] = tp
            os.environ["LVB_GPU_PIECE_MIN"] = pm
            os.environ["LVB_GPU_STORE_STREAM"] = st
            os.environ["LVB_GPU_STORE_PF"] = pf if int(st) == 0 else "0"
            os.environ["LVB_GPU_STREAM_LAG"] = pf if int(st) > 0 else "-1"
            os.environ["LVB_GPU_TARGET_WARPS"] = str(148 * int(tw))
            os.environ["LVB_GPU_BLOCK"] = blk
            eng = Engine(0)
            try:
                eng.upload(enc, word_begin=0, word_count=wc, max_trees=1)
                tree = eng.tree_create()
                length = eng.commit(tree, topo)
                for _ in range(5):
                    eng.rerun_last_batch()
                eng.synchronize()
                ext = torch.cuda.ExternalStream(eng.stream)
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                eng.kernel_time(reset=True)
                l0 = eng.launches
                e0.record(ext)
                for _ in range(args.reps):
                    eng.rerun_last_batch()
                e1.record(ext)
                eng.synchronize()
                k_ms, k_n = eng.kernel_time()
                key = shards
                want.setdefault(key, length)
                print(json.dumps({"shards": shards, "words": wc, "stream": int(st), "taper": int(tp), "piece_min": int(pm), "pf_or_lag": int(pf), "target_warps_per_sm": int(tw),
                                  "block": int(blk), "kernel_ms": round(k_ms / max(1, k_n), 4),
                                  "stream_ms": round(e0.elapsed_time(e1) / args.reps, 4),
                                  "launches_per_pass": (eng.launches - l0) / args.reps, "length": int(length),
                                  "length_ok": bool(length == want[key])}), flush=True)
            except Exception as exc:
                print(json.dumps({"shards": shards, "stream": int(st), "pf_or_lag": int(pf), "target_warps_per_sm": int(tw), "block": int(blk),
                                  "error": str(exc)[:200]}), flush=True)
            finally:
                eng.close()


if __name__ == "__main__":
    main()
