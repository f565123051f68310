#!/usr/bin/env python
"""bench.py — Fitch node x site updates/s and SA trees scored/s (BASELINE.json metric) on B200.

Workload (config.workload): BASELINE.json configs[3] shape, 2000 taxa x 1,000,000 sites.  A step is
one full getplen pass (every internal node dirty: n-3 stored Fitch combinations + the 2 root-step
applications = (n-1) * m node x site updates) of ONE tree of that shape.

N = 1
  value  device-resident: the program of the tree is already on the device, the step is the
         launch of the Fitch kernel (counts are folded inside it), CUDA events on the engine's stream.
  e2e    through the C-ABI (lvb_gpu_tree_commit) with HOST topology arrays: host schedule compile,
         H2D of the program, kernel, D2H of the tree length, every step, each step a different tree.
N > 1 (the driver's torchrun command, one rank per GPU) — the two partitions north_star names:
  value  BASELINE configs[3]: the SAME one tree, its words sharded over the ranks in site blocks
         (the reference's OpenMP decomposition, TreeEvaluation.c:95-97), partial lengths summed by
         the all-reduce inside the C-ABI; "scaling": "strong".  The all-reduced length must equal
         the single-GPU length of the same tree (asserted).
  replicas       sub-record: one whole tree per GPU per step, no collective (weak scaling).
  sa_replicates  sub-record, BASELINE configs[4]: 64 SA chains (seeds 1000+r) on 1000 x 200k through
         the C++ SA driver (lvb_b200/host/solve.cpp), chains sharded over the ranks
         (sharding.replicate_shard), fixed rearrangement budget per chain; trees scored/s and
         rearrangements/s summed over ranks.  Printed only if sampled result trees re-evaluate to the
         reported lengths on the CPU oracle in the same run.

--impl reference times the reference's own CPU getplen (oracle/_ref, OpenMP, all host cores) on the
same 2000 x 1M alignment and tree.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import tempfile
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "fitch_node_site_updates_per_s"
UNIT = "node*site updates/s"


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--mode", default="auto", choices=["auto", "replicas", "siteblock"],
                    help="headline at N > 1: auto/siteblock = one tree in site blocks + all-reduce; replicas = one tree per GPU")
    ap.add_argument("--taxa", type=int, default=2000)
    ap.add_argument("--sites", type=int, default=1_000_000)
    ap.add_argument("--cpu-sample-sites", type=int, default=0,
                    help="CPU legs on a column sample of this many sites (0 = the full alignment; a sample is only "
                         "taken by itself when host memory is short)")
    ap.add_argument("--candidates", type=int, default=512)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--sa-chains", type=int, default=64)
    ap.add_argument("--sa-taxa", type=int, default=1000)
    ap.add_argument("--sa-sites", type=int, default=200_000)
    ap.add_argument("--sa-rearrangements", type=int, default=3000)
    ap.add_argument("--no-sa", action="store_true")
    ap.add_argument("--no-extras", action="store_true", help="headline only (no candidate scoring / NNI / SA sub-records)")
    return ap.parse_args()


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md)"


def traffic_from_profile(key="dram_bytes_per_pass"):
    """DRAM bytes per launch of a kernel from the committed `ncu --set full` capture
    (profiles/ncu_traffic.json, written by tools/ncu_summary.py); None if there is none."""
    p = os.path.join(ROOT, "profiles", "ncu_traffic.json")
    if not os.path.exists(p):
        return None, None
    d = json.load(open(p))
    return d.get(key), d.get("source")


def host_memory_gb() -> float:
    try:
        for line in open("/proc/meminfo"):
            if line.startswith("MemAvailable:"):
                return int(line.split()[1]) / 1e6
    except OSError:
        pass
    return 0.0


class ClockSampler:
    """SM clock and throttle reasons sampled DURING the timed regions only (NVML, every 4 ms while
    `active` is set — the device idles between the regions, and idle clocks say nothing about the run;
    nvidia-smi as a fallback)."""

    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index, self.samples, self.stop = index, [], threading.Event()
        self.active = False
        self.thread = threading.Thread(target=self.run, daemon=True)
        self.nvml = None
        try:
            import pynvml

            pynvml.nvmlInit()
            # NVML indexes physical devices; honour CUDA_VISIBLE_DEVICES when it lists indices
            vis = os.environ.get("CUDA_VISIBLE_DEVICES", "")
            phys = index
            if vis and all(x.strip().isdigit() for x in vis.split(",")):
                phys = int(vis.split(",")[index])
            self.handle = pynvml.nvmlDeviceGetHandleByIndex(phys)
            self.max_sm = pynvml.nvmlDeviceGetMaxClockInfo(self.handle, pynvml.NVML_CLOCK_SM)
            self.nvml = pynvml
        except Exception:
            self.nvml = None

    def run(self):
        if self.nvml is not None:
            nv = self.nvml
            bits = [("hw_slowdown", nv.nvmlClocksThrottleReasonHwSlowdown),
                    ("hw_thermal_slowdown", nv.nvmlClocksThrottleReasonHwThermalSlowdown),
                    ("sw_thermal_slowdown", nv.nvmlClocksThrottleReasonSwThermalSlowdown),
                    ("sw_power_cap", nv.nvmlClocksThrottleReasonSwPowerCap)]
            while not self.stop.is_set():
                if not self.active:
                    self.stop.wait(0.001)
                    continue
                try:
                    sm = nv.nvmlDeviceGetClockInfo(self.handle, nv.NVML_CLOCK_SM)
                    r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.handle)
                    self.samples.append([str(sm), str(self.max_sm)] + ["Active" if r & b else "Not Active" for _, b in bits])
                except Exception:
                    pass
                self.stop.wait(0.004)
            return
        while not self.stop.is_set():
            if not self.active:
                self.stop.wait(0.001)
                continue
            try:
                out = subprocess.run(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-i",
                                      str(self.index)], capture_output=True, text=True, timeout=5).stdout.strip()
                if out:
                    self.samples.append([x.strip() for x in out.split(",")])
            except Exception:
                pass
            self.stop.wait(0.1)

    def __enter__(self):
        self.thread.start()
        return self

    def __exit__(self, *a):
        self.stop.set()
        self.thread.join(timeout=6)

    def summary(self):
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no clock samples"]}
        sm = sorted(int(s[0]) for s in self.samples if s[0].isdigit())
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [nm for k, nm in enumerate(names) if any(s[2 + k].lower().startswith("active") for s in self.samples)]
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": int(self.samples[0][1]),
                "reasons": reasons, "samples": len(self.samples), "source": "nvml" if self.nvml else "nvidia-smi"}


def parent_of(left, right, root, n):
    B = 2 * n - 3
    parent = np.full(B, -1, dtype=np.int64)
    for v in list(range(n, B)) + [root]:
        parent[left[v]] = v
        parent[right[v]] = v
    return parent


def cpu_reference_leg(enc, m, left, right, root, threads, passes, warm=1):
    """The reference's own getplen (oracle/_ref, unmodified, OpenMP) on `enc`: `passes` timed full
    passes of the tree after `warm` untimed ones.  Returns None if the reference was not built."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import oracles

    if not oracles.have_ref():
        return None
    import ctypes as C

    n = enc.shape[0]
    msa = oracles.RefMSA.from_packed(enc, m, threads=threads)
    t = msa.tree()
    t.set_topology(parent_of(left, right, root, n), left, right, root)
    L = oracles.ref()
    length = C.c_long(0)
    if warm > 0:
        L.ref_time_full_passes(msa.h, t.t, root, warm, C.byref(length))
    secs = L.ref_time_full_passes(msa.h, t.t, root, passes, C.byref(length))
    updates = (n - 1) * msa.m * passes
    return {
        "value": updates / secs, "unit": UNIT, "cores": int(msa.n_threads), "kind": "reference",
        "sample": f"{passes} full getplen passes (unmodified reference, OpenMP team of {msa.n_threads}, slices of "
                  f"{msa.slice} words) of the same tree on {n} taxa x {msa.m} sites; {secs / passes * 1e3:.1f} ms/pass",
        "ms_per_pass": secs / passes * 1e3, "length": int(length.value), "sites": int(msa.m),
    }


def cpu_workload(args, n, m):
    """(enc, sites) the CPU legs run on: the full alignment unless memory is short or a sample was asked for"""
    from lvb_b200 import synthetic

    sites = m
    if args.cpu_sample_sites > 0:
        sites = min(m, args.cpu_sample_sites)
    elif host_memory_gb() < 12.0 and m > 100_000:
        sites = 100_000  # 2000 x 1M needs 1 GB packed + 2 GB of reference tree rows (+ this process)
    return sites


def nni_candidates(left, right, root, n, count, seed):
    """`count` NNI rearrangements of the tree with their dirty paths (host test-input code)"""
    from lvb_b200 import Topology

    B = 2 * n - 3
    parent = parent_of(left, right, root, n)
    rng = np.random.default_rng(seed)
    out, index = [], []  # index: position of the candidate in lvb_gpu_nni_neighbourhood's output
    for u in rng.integers(n, B, size=count).tolist():
        v = int(parent[u])
        a, b = int(left[u]), int(right[u])
        c = int(right[v]) if left[v] == u else int(left[v])
        l2, r2 = left.copy(), right.copy()
        swap_a = bool(rng.integers(0, 2))
        moved = a if swap_a else b
        if l2[v] == u:
            r2[v] = moved
        else:
            l2[v] = moved
        l2[u], r2[u] = (b, c) if swap_a else (a, c)
        dirty, w = [], u
        while w != root:
            dirty.append(w)
            w = int(parent[w])
        # the candidate as LVB's operators leave it for the device layer: the current tree's arrays plus the two
        # nodes whose links the rearrangement rewired (lvb_gpu_topology::mod_*), not a private copy of 2 x 32 kB
        cand = Topology(left, right, root, dirty, overlay=([u, v], [l2[u], l2[v]], [r2[u], r2[v]]))
        cand.full = (l2, r2)
        out.append(cand)
        index.append(2 * (u - n) + (1 if swap_a else 0))
    return out, index


# ---------------------------------------------------------------------------------------------
# SA replicates (BASELINE configs[4]) through the C++ SA driver
# ---------------------------------------------------------------------------------------------
def newick_topologies(text: str, titles):
    """trees of an `outtree` file as LVB arrays (left, right, root): lvb_treeprint writes
    (rootleaf,subtree,subtree); — src/TreeOperations.c:1144-1220"""
    index = {t: i for i, t in enumerate(titles)}
    n = len(titles)
    out = []
    for line in text.split(";"):
        s = line.strip()
        if not s:
            continue
        left = np.full(2 * n - 3, -1, dtype=np.int64)
        right = np.full(2 * n - 3, -1, dtype=np.int64)
        nxt = [n]
        pos = [0]

        def parse():
            if s[pos[0]] == "(":
                pos[0] += 1
                kids = [parse()]
                while s[pos[0]] == ",":
                    pos[0] += 1
                    kids.append(parse())
                assert s[pos[0]] == ")"
                pos[0] += 1
                return kids
            j = pos[0]
            while s[j] not in ",()":
                j += 1
            name = s[pos[0]:j].strip()
            pos[0] = j
            return index[name]

        def build(node):
            if not isinstance(node, list):
                return node
            assert len(node) == 2
            a, b = build(node[0]), build(node[1])
            v = nxt[0]
            nxt[0] += 1
            left[v], right[v] = a, b
            return v

        top = parse()
        assert isinstance(top, list) and len(top) == 3 and not isinstance(top[0], list)
        root = top[0]
        left[root], right[root] = build(top[1]), build(top[2])
        assert nxt[0] == 2 * n - 3
        out.append((left, right, int(root)))
    return out


class c_stdout_to_stderr:
    """the C++ search prints LVB's progress table with printf; bench.py's stdout carries ONE JSON line"""

    def __enter__(self):
        sys.stdout.flush()
        self.saved = os.dup(1)
        os.dup2(2, 1)

    def __exit__(self, *a):
        import ctypes

        ctypes.CDLL(None).fflush(None)
        os.dup2(self.saved, 1)
        os.close(self.saved)


def sa_replicates_record(args, rank, world, local_rank, barrier, gather):
    """64 chains of 1000 x 200k sharded over the ranks; returns the sub-record on rank 0 (None elsewhere)"""
    from lvb_b200 import host, sharding, synthetic

    n, m, R = args.sa_taxa, args.sa_sites, args.sa_chains
    st = synthetic.survey_alignment(n, m, 1, 0.02)
    rows = synthetic.text_rows(st)
    msa = host.HostMSA.from_rows(rows)
    seed0, count = sharding.replicate_shard(1000, R, rank, world)
    threads = max(1, (os.cpu_count() or 1) // max(1, min(world, 8)))
    with tempfile.TemporaryDirectory() as td, c_stdout_to_stderr():
        outtree = os.path.join(td, "outtree")
        # warm-up: CUDA context, kernels, pools (a short search of the same shape)
        msa.search(seed0, algorithm=1, device=local_rank, replicates=min(count, 4), max_rearrangements=100,
                   host_threads=threads)
        # the search is host-driven and latency-bound: a busy host core shows at once (0.56 - 0.81 s for the same job on
        # boxes of one pool).  Three complete runs of the same job, every one reported, the fastest one quoted.
        runs = []
        for _ in range(3):
            barrier()
            t0 = time.perf_counter()
            res_i = msa.search(seed0, algorithm=1, device=local_rank, replicates=count, max_rearrangements=args.sa_rearrangements,
                               host_threads=threads, outtree_path=outtree)
            wall_i = time.perf_counter() - t0
            barrier()
            runs.append((res_i[0]["seconds"], wall_i, res_i))
        all_search_s = [r[0] for r in runs]
        all_wall_s = [r[1] for r in runs]
        res = runs[-1][2]  # chains are deterministic: every run gives the same trees and counts
        # parity gate of this very run: the best tree of the rank's first and last chain, re-evaluated by the CPU oracle
        sys.path.insert(0, os.path.join(ROOT, "tests"))
        import oracles

        enc = msa.enc()
        titles = [msa.title(i) for i in range(msa.n)]
        checked, ok = [], True
        for r in sorted({0, count - 1}):
            path = outtree if r == 0 else f"{outtree}.r{r}"
            left, right, root = newick_topologies(open(path).read(), titles)[0]
            want = oracles.oracle_getplen(enc, left, right, root)[0]
            checked.append({"chain": seed0 - 1000 + r, "reported": res[r]["final_length"], "oracle": want})
            ok = ok and want == res[r]["final_length"]
    mine = {"rank": rank, "chains": count, "host_threads": threads,
            "search_s_all_runs": all_search_s, "wall_s_all_runs": all_wall_s,
            "trees_scored": sum(r["trees_scored"] for r in res), "rearrangements": sum(r["rearrangements"] for r in res),
            "groups": res[0]["groups"], "checked": checked, "ok": ok,
            "final_lengths_head": [r["final_length"] for r in res[:2]]}
    allr = gather(mine)
    msa.close()
    if rank != 0:
        return None
    if not all(r["ok"] for r in allr):
        return {"refused": "a sampled result tree did not re-evaluate to the reported length on the oracle",
                "checked": [c for r in allr for c in r["checked"]]}
    # the SA driver's own clock (GetSolnBatched: opening the driver groups, first evaluation of every start tree, the
    # whole search), max over ranks; wall_s adds the one-time set-up of the call (DNAToBinary on the device, leaf rows,
    # 0.1 GB of state rows per chain allocated) and writing the result trees
    # the run (same index on every rank: the runs are separated by barriers) whose slowest rank was fastest
    n_runs = len(allr[0]["search_s_all_runs"])
    best = min(range(n_runs), key=lambda i: max(r["search_s_all_runs"][i] for r in allr))
    for r in allr:
        r["search_s"], r["wall_s"] = r["search_s_all_runs"][best], r["wall_s_all_runs"][best]
        r["search_s_all_runs"] = [round(x, 4) for x in r["search_s_all_runs"]]
    secs = max(r["search_s"] for r in allr)
    wall = max(r["wall_s"] for r in allr)
    return {
        "workload": f"{R} SA chains (seeds 1000..{1000 + R - 1}), {n} taxa x {m} sites, {args.sa_rearrangements} rearrangements "
                    f"per chain, algorithm 1; chains sharded over {world} rank(s), one process per GPU, no collective",
        "trees_scored_per_s": sum(r["trees_scored"] for r in allr) / secs,
        "rearrangements_per_s": sum(r["rearrangements"] for r in allr) / secs,
        "seconds": secs, "wall_seconds": wall,
        "trees_scored_per_s_wall": sum(r["trees_scored"] for r in allr) / wall,
        "timing": "seconds = the SA driver's own clock around the search (host-driven loop, set-up of its driver groups "
                  "included), max over ranks; wall_seconds = wall clock of the whole call between barriers (adds packing "
                  "and uploading the alignment, allocating the trees, writing the result files)",
        "per_rank": [{k: r[k] for k in ("rank", "chains", "search_s", "search_s_all_runs", "wall_s", "host_threads", "groups")}
                     for r in allr],
        "runs": "3 complete runs of the job, all ranks together each time; quoted is the run whose slowest rank was "
                "fastest (search_s_all_runs has all of them)",
        "scaling_note": "strong scaling of a FIXED job (64 chains x the rearrangement budget): a chain is a sequence of "
                        "rounds (one per accepted move), each waiting for one latency-bound kernel, so the job cannot finish "
                        "before one chain alone does (about 45 us per round, profiles/r02n_sa_probe.log); with 64 chains "
                        "on one GPU the device is throughput-bound instead, so trees/s at N GPUs is bounded by about 4x the "
                        "N = 1 figure however many GPUs there are; ranks share the host cores (host_threads per rank)",
        "parity": {"oracle_checked": [c for r in allr for c in r["checked"]], "ok": True},
        "api": "lvb_b200/host/solve.cpp GetSolnBatched (speculative batches over lvb_gpu_score_batch_submit_pre / "
               "lvb_gpu_commit_batch_pre)",
    }


def main():
    args = parse_args()
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    n, m = args.taxa, args.sites
    nwords = (m + 15) // 16
    nproc = os.cpu_count() or 1

    from lvb_b200 import synthetic

    if args.impl == "reference":
        if rank != 0:
            return
        t0 = time.time()
        sites = cpu_workload(args, n, m)
        enc = synthetic.random_packed(n, nwords, seed=1)
        if sites != m:
            enc = np.ascontiguousarray(enc[:, : (sites + 15) // 16])
        left, right, root = synthetic.random_topology(n, seed=3)
        # every step is one full pass of the tree; W warm-up passes, then K timed ones (bounded: the whole run
        # stays within minutes at ~0.1-0.2 s per pass)
        passes = max(1, min(args.steps, 60))
        leg = cpu_reference_leg(enc, sites, left, right, root, nproc, passes, warm=max(1, min(args.warmup, 5)))
        if leg is None:
            print(json.dumps({"impl": "reference", "unavailable": "oracle/_ref not built on this box"}))
            return
        line = {
            "impl": "reference", "metric": METRIC, "value": leg["value"], "unit": UNIT, "n_gpus": args.gpus,
            "steps": passes, "warmup": max(1, min(args.warmup, 5)), "ms_per_step": leg["ms_per_pass"],
            "higher_is_better": True, "scaling": "strong" if args.gpus > 1 else "weak", "vs_baseline": None,
            "dtype": "u64 (4-bit state sets, 64-bit SWAR words)", "data": "synthetic",
            "config": {"workload": f"full getplen pass, {n} taxa x {leg['sites']} sites"
                                   + ("" if leg["sites"] == m else f" (column sample of the {n} x {m} workload)")
                                   + f", one tree, reference CPU build, {leg['cores']} OpenMP threads",
                       "taxa": n, "sites": leg["sites"], "same_config_as_b200_arm": leg["sites"] == m},
            "cpu_baseline": {k: leg[k] for k in ("value", "unit", "cores", "kind", "sample")},
            "e2e": {"value": leg["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "length_check": leg["length"], "wall_s": time.time() - t0,
        }
        print(json.dumps(line))
        return

    import torch

    dist = None
    if world > 1:
        import torch.distributed as dist

        torch.cuda.set_device(local_rank)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x: float) -> float:
        if dist is None:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def gather(obj):
        if dist is None:
            return [obj]
        out = [None] * world
        dist.all_gather_object(out, obj)
        return out

    from lvb_b200 import Engine, Topology

    siteblock = world > 1 and args.mode in ("auto", "siteblock")
    enc = synthetic.random_packed(n, nwords, seed=1)
    left, right, root = synthetic.random_topology(n, seed=3)
    topo = Topology(left, right, root)
    updates_per_pass = (n - 1) * m  # real sites; the reference's count
    peak, peak_src = peaks()

    def timed_passes(eng, steps):
        """`steps` device-resident passes of the batch resident on eng: (seconds by CUDA events on the engine's
        stream [max over ranks], this rank's mean kernel ms, launches)"""
        ext = torch.cuda.ExternalStream(eng.stream, device=torch.device("cuda", local_rank))
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        l0 = eng.launches
        eng.kernel_time(reset=True)
        barrier()
        clocks.active = True
        ev0.record(ext)
        for _ in range(steps):
            eng.rerun_last_batch()
        ev1.record(ext)
        eng.synchronize()
        clocks.active = False
        barrier()
        mine = ev0.elapsed_time(ev1) * 1e-3
        k_ms, k_launch = eng.kernel_time()
        return max_over_ranks(mine), mine, k_ms / max(1, k_launch), eng.launches - l0

    warm = max(3, args.warmup)
    extras = {}
    with ClockSampler(local_rank) as clocks:
        # =================== headline ===================
        eng = Engine(local_rank)
        if siteblock:
            from lvb_b200 import sharding

            w0, wc = sharding.word_block(nwords, rank, world)
            eng.upload(enc, word_begin=w0, word_count=wc, max_trees=2)
            uid = [Engine.comm_unique_id() if rank == 0 else None]
            dist.broadcast_object_list(uid, src=0)
            eng.comm_init(world, rank, uid[0])
            tree_seed = 3
        else:
            eng.upload(enc, max_trees=2)
            tree_seed = 3 + 1000 * rank  # replicas: every rank its own trees
            if rank:
                left, right, root = synthetic.random_topology(n, seed=tree_seed)
                topo = Topology(left, right, root)
        my_words = eng.word_count
        tree = eng.tree_create()
        length0 = eng.commit(tree, topo)
        for _ in range(warm):
            eng.rerun_last_batch()
        eng.synchronize()
        # ---- region A: device-resident ----
        wall_a, wall_a_mine, kernel_ms_mine, gpu_launches = timed_passes(eng, args.steps)
        kernel_ms = max_over_ranks(kernel_ms_mine)
        per_rank = gather({"rank": rank, "kernel_ms": kernel_ms_mine, "stream_ms_per_step": wall_a_mine / args.steps * 1e3,
                           "words": int(my_words), "length": int(length0)})
        if siteblock:  # every rank got the all-reduced length of the one tree
            assert all(r["length"] == per_rank[0]["length"] for r in per_rank), per_rank

        # ---- region B: end to end through the C-ABI, host topology arrays in, length out ----
        e2e_seed = 3 if siteblock else tree_seed
        topos = [Topology(*synthetic.random_topology(n, seed=e2e_seed + 17 * (s + 1))) for s in range(args.steps)]
        for s in range(min(3, args.steps)):
            eng.commit(tree, topos[s])
        barrier()
        clocks.active = True
        t0 = time.perf_counter()
        # what a caller that scores many trees does: tree s+1 is compiled and enqueued (lvb_gpu_commit_batch_submit)
        # while the device works on tree s; every length is read back (lvb_gpu_batch_wait)
        lens, inflight = [], []
        for tp in topos:
            inflight.append(eng.commit_batch_submit([tree], [tp]))
            if len(inflight) > 2:
                lens.append(int(eng.batch_wait(inflight.pop(0))[0]))
        while inflight:
            lens.append(int(eng.batch_wait(inflight.pop(0))[0]))
        clocks.active = False
        barrier()
        wall_b = max_over_ranks(time.perf_counter() - t0)
        # the same, one tree at a time (submit, wait, next): the latency of a single call
        t0 = time.perf_counter()
        for tp in topos[: max(1, args.steps // 4)]:
            eng.commit(tree, tp)
        barrier()
        wall_b_sync = max_over_ranks(time.perf_counter() - t0) / max(1, args.steps // 4)

        total_ranks = 1 if siteblock else world
        value = total_ranks * args.steps * updates_per_pass / wall_a
        e2e_value = total_ranks * args.steps * updates_per_pass / wall_b

        if not args.no_extras and world == 1:
            # ---- SA candidate scoring: trees scored/s ----
            eng.commit(tree, topo)
            cands, cand_index = nni_candidates(left, right, root, n, args.candidates, seed=5)
            trees = [tree] * len(cands)
            got = eng.score_batch(trees, cands)
            prepared = eng.prepare_batch(trees, cands)  # the C arrays a C caller holds anyway
            eng.score_prepared(prepared)
            reps_c, calls = 9, []
            for _ in range(reps_c):
                t0 = time.perf_counter()
                got = eng.score_prepared(prepared).copy()
                calls.append(time.perf_counter() - t0)
            wall_c = float(np.median(calls)) * reps_c  # per-call median: one host hiccup in nine calls is not the engine
            t0 = time.perf_counter()
            eng.score_batch(trees, cands)
            wall_py = time.perf_counter() - t0
            eng.kernel_time(reset=True)
            for _ in range(3):
                eng.rerun_last_batch()
            eng.synchronize()
            c_ms, c_launch = eng.kernel_time()
            dirty_mean = float(np.mean([len(c.dirty) for c in cands]))
            from lvb_b200 import compile_schedule

            rows_read = [int((compile_schedule(n, c, store=False)[0][:, :2] >= 0).sum()) for c in cands[:64]]
            rows_read_mean = float(np.mean(rows_read))
            score_ms = c_ms / max(1, c_launch)
            score_traffic, _ = traffic_from_profile("score_dram_bytes_per_batch")
            prof = json.load(open(os.path.join(ROOT, "profiles", "ncu_traffic.json"))) if score_traffic else {}
            extras["sa_trees_scored_per_s"] = {
                "e2e": reps_c * len(cands) / wall_c, "device": len(cands) / (score_ms * 1e-3),
                "batch": len(cands), "mean_dirty_nodes": dirty_mean,
                "e2e_through_python_binding": len(cands) / wall_py,
                "e2e_ms_per_call": {"median": float(np.median(calls)) * 1e3, "min": min(calls) * 1e3, "max": max(calls) * 1e3},
                "api": "lvb_gpu_score_batch: NNI candidates of the current tree, one launch per batch; e2e = the C-ABI call "
                       "with host topology arrays (schedule compile on the host threads, H2D of the programs, kernel, lengths "
                       "back), the lvb_gpu_topology records prepared once; e2e_through_python_binding adds building those "
                       "records in Python every call; the C++ SA driver is measured in sa_replicates",
            }
            operand_gbs = rows_read_mean * 8 * my_words * len(cands) / (score_ms * 1e-3) / 1e9
            extras["roofline_score"] = {
                "bound": "integer issue (ALU pipe), not HBM: candidates of one tree share their operand rows through L2",
                "unit": "GB/s", "peak": peak,
                "achieved": (score_traffic / (score_ms * 1e-3) / 1e9) if score_traffic else None,
                "frac": (score_traffic / (score_ms * 1e-3) / 1e9 / peak) if score_traffic else None,
                "traffic": score_traffic, "traffic_source": prof.get("score_source"),
                "l2_hit_pct": prof.get("score_l2_hit_pct"), "alu_pipe_pct": prof.get("score_alu_pipe_pct"),
                "issue_pct": prof.get("score_issue_pct"),
                "operand_gbs": operand_gbs, "operand_frac": operand_gbs / peak,
                "rows_read_per_candidate": rows_read_mean, "kernel": prof.get("score_kernel", "fitch_score2_kernel"),
                "kernel_ms_per_batch": score_ms,
                "note": "achieved/frac = DRAM bytes per batch of the committed ncu capture / this run's kernel time / HBM peak "
                        "(low by design: 80 % of the operand sectors hit L2); what bounds the kernel is instruction issue, "
                        "alu_pipe_pct / issue_pct of the same capture; operand_* = 8 B x nwords x (clean rows read) per "
                        "candidate (SURVEY 8(d)), served by L2 + HBM together",
            }
            # ---- exhaustive NNI neighbourhood (deterministic_hillclimb's 2(n-3) candidates) in one call ----
            scratch = eng.tree_create()
            nn = eng.nni_neighbourhood(tree, scratch, topo)
            t0 = time.perf_counter()
            reps_n = 3
            for _ in range(reps_n):
                nn = eng.nni_neighbourhood(tree, scratch, topo)
            wall_n = time.perf_counter() - t0
            nn_ok = bool(np.array_equal(nn[np.asarray(cand_index)], np.asarray(got)))
            extras["nni_neighbourhood"] = {
                "candidates": int(len(nn)), "ms_per_call": wall_n / reps_n * 1e3,
                "candidates_per_s": len(nn) / (wall_n / reps_n), "four_row_kernel_ms": eng.last_kernel_ms(),
                "agrees_with_score_batch": nn_ok,
                "api": "lvb_gpu_nni_neighbourhood: up-pass program + four-row kernel, host arrays in, 2(n-3) lengths out",
            }
        eng.close()

        # =================== N > 1: the weak-scaling sub-record (one whole tree per GPU) ===================
        if siteblock and not args.no_extras:
            eng2 = Engine(local_rank)
            eng2.upload(enc, max_trees=1)
            t2 = eng2.tree_create()
            l2, r2, root2 = synthetic.random_topology(n, seed=3 + 1000 * rank)
            eng2.commit(t2, Topology(l2, r2, root2))
            for _ in range(warm):
                eng2.rerun_last_batch()
            eng2.synchronize()
            w_all, w_mine, k_mine, _ = timed_passes(eng2, args.steps)
            rr = gather({"rank": rank, "kernel_ms": k_mine, "stream_ms_per_step": w_mine / args.steps * 1e3})
            extras["replicas"] = {
                "value": world * args.steps * updates_per_pass / w_all, "unit": UNIT, "scaling": "weak",
                "ms_per_step": w_all / args.steps * 1e3, "per_rank": rr,
                "workload": f"one whole {n} x {m} tree per GPU per step, alignment replicated, no collective",
            }
            # ---- candidates of ONE tree sharded over the GPUs (third row): every rank holds the whole
            # alignment and the same tree and scores its range of the batch; the lengths are gathered by the host
            from lvb_b200 import sharding as _sh

            eng2.commit(t2, topo)
            cands, _ = nni_candidates(left, right, root, n, args.candidates, seed=5)
            first, count = _sh.candidate_range(len(cands), rank, world)
            mine_c = cands[first:first + count]
            prepared = eng2.prepare_batch([t2] * count, mine_c)
            eng2.score_prepared(prepared)
            calls = []
            for _ in range(5):
                barrier()
                t0 = time.perf_counter()
                got_c = eng2.score_prepared(prepared).copy()
                calls.append(max_over_ranks(time.perf_counter() - t0))
            kms = eng2.last_kernel_ms()
            parts = gather({"rank": rank, "first": first, "lengths": got_c.tolist(), "kernel_ms": kms})
            # cross-check: the first candidate of every other rank's range, scored here too
            probe = [cands[r["first"]] for r in parts if r["lengths"]]
            probe_len = eng2.score_batch([t2] * len(probe), probe).tolist()
            ok = probe_len == [r["lengths"][0] for r in parts if r["lengths"]]
            extras["candidate_shards"] = {
                "candidates": len(cands), "ms_per_batch": float(np.median(calls)) * 1e3,
                "candidates_per_s": len(cands) / float(np.median(calls)), "kernel_ms_per_rank": [r["kernel_ms"] for r in parts],
                "cross_checked": bool(ok),
                "workload": f"{len(cands)} NNI candidates of one {n} x {m} tree, contiguous ranges over the ranks "
                            "(sharding.candidate_range), whole alignment and the same tree on every GPU, lengths gathered on "
                            "the host; the C-ABI call with host topology records, slowest rank, median of 5",
            }
            assert ok, "candidate shards: a rank's lengths differ from another rank's for the same candidate"
            eng2.close()

    # =================== SA replicates (config 5) ===================
    if not args.no_sa and not args.no_extras:
        try:
            rec = sa_replicates_record(args, rank, world, local_rank, barrier, gather)
        except Exception as exc:  # a sub-record never takes the headline down
            rec = {"failed": f"{type(exc).__name__}: {exc}"[:300]}
        if rank == 0:
            extras["sa_replicates"] = rec

    # =================== the line ===================
    algo_bytes_per_pass = (3 * n - 6) * 8 * my_words  # 24 B per node-word update, SURVEY.md §8(d)
    fused_bytes_per_pass = (2 * n - 3) * 8 * my_words  # every leaf row read once, every internal row written once
    traffic, traffic_src = traffic_from_profile()
    if traffic is not None and world > 1:
        traffic = traffic * my_words / nwords  # the capture is of the whole-width pass
    phys = traffic if traffic is not None else fused_bytes_per_pass
    achieved = phys / (kernel_ms * 1e-3) / 1e9
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": warm,
        "ms_per_step": wall_a / args.steps * 1e3, "higher_is_better": True,
        "scaling": "strong" if siteblock else "weak", "vs_baseline": None, "dtype": "u64 (4-bit state sets, 32-bit lanes)",
        "data": "synthetic",
        "config": {
            "workload": f"full getplen pass, {n} taxa x {m} sites (BASELINE.json configs[3] shape), "
                        + ("ONE tree, words of every row in site blocks over the ranks, partial lengths all-reduced "
                           "inside the C-ABI" if siteblock else "one tree per step per GPU, alignment resident on the GPU"),
            "taxa": n, "sites": m, "words_per_row": nwords, "state_bytes_per_tree": (2 * n - 3) * 8 * nwords,
            "l2": "inputs larger than L2: 1.0 GB alignment + 2.0 GB state rows per pass vs 126 MB L2"
                  + ("" if not siteblock else f" ({3.0 / world:.2f} GB per rank)"),
            "parallelism": ("siteblock" if siteblock else "replicas") + str(world),
        },
        "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": (n - 1) * 16 + 64, "d2h_bytes_per_step": 8,
                "ms_per_step": wall_b / args.steps * 1e3, "ms_per_call_unpipelined": wall_b_sync * 1e3,
                "api": "lvb_gpu_commit_batch_submit(host left/right/root) + lvb_gpu_batch_wait(-> length), a different "
                       "random tree every step, up to 3 calls in flight (host compile of tree s+1 under the kernel of "
                       "tree s); ms_per_call_unpipelined = lvb_gpu_tree_commit one call at a time"},
        "gpu_launches": gpu_launches,
        "per_rank": per_rank,
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                     "traffic": traffic, "traffic_source": traffic_src if traffic is not None else None,
                     "traffic_unit": "DRAM bytes per pass (ncu dram__bytes_read.sum + dram__bytes_write.sum)",
                     "bytes_used": phys,
                     "bytes_used_is": "measured DRAM traffic (ncu)" if traffic is not None else
                                      "floor of the fused algorithm: (2n-3) rows x 8 B x words",
                     "fused_floor_bytes_per_pass": fused_bytes_per_pass,
                     "frac_on_fused_floor": fused_bytes_per_pass / (kernel_ms * 1e-3) / 1e9 / peak,
                     "algorithmic_bytes_per_pass": algo_bytes_per_pass,
                     "algorithmic_gbs": algo_bytes_per_pass / (kernel_ms * 1e-3) / 1e9,
                     "algorithmic_frac": algo_bytes_per_pass / (kernel_ms * 1e-3) / 1e9 / peak,
                     "peak_source": peak_src, "kernel": "the store kernel of the pass (one launch per pass)",
                     "timing": "CUDA events on the engine's stream around each launch, mean over the timed steps, max over ranks",
                     "kernel_ms_per_pass": kernel_ms,
                     "note": "frac = physical DRAM bytes / kernel time / peak.  SURVEY 8(d)'s 24 B per node-word (2 reads + "
                             "1 write) is kept as algorithmic_*: the kernel keeps interior operands on chip and moves ~16 B "
                             "per node-word, so algorithmic_frac can exceed 1"},
        "clocks": clocks.summary(),
        "length_check": int(length0), "e2e_lengths_head": [int(x) for x in lens[:3]],
    }
    line.update(extras)
    if siteblock:
        line["length_expected_single_gpu"] = 1101822927 if (n, m) == (2000, 1_000_000) else None
        if line["length_expected_single_gpu"] is not None:
            assert int(length0) == line["length_expected_single_gpu"], (length0, line["length_expected_single_gpu"])
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        try:
            sites = cpu_workload(args, n, m)
            cenc = enc if sites == m else np.ascontiguousarray(enc[:, : (sites + 15) // 16])
            leg = cpu_reference_leg(cenc, sites, left, right, root, nproc, passes=8 if sites == m else 40)
            if leg is not None:
                line["cpu_baseline"] = {k: leg[k] for k in ("value", "unit", "cores", "kind", "sample")}
                if sites == m:
                    line["length_match"] = bool(leg["length"] == int(length0))
                    line["cpu_baseline"]["length"] = leg["length"]
        except Exception as exc:  # the baseline is reported, never required
            line["cpu_baseline"] = {"value": None, "unit": UNIT, "cores": nproc, "kind": "reference",
                                    "sample": f"failed: {exc}"}
    if rank == 0:
        print(json.dumps(line))
    if dist is not None:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
