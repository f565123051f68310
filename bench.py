#!/usr/bin/env python
"""bench.py — Fitch node x site updates/s (BASELINE.json metric) on B200.

Workload (config.workload): BASELINE.json configs[3] shape, 2000 taxa x 1,000,000 sites,
which fits one GPU (1.0 GB alignment + 2.0 GB of state rows).  A step is one full getplen
pass (every internal node dirty: n-3 stored Fitch combinations + the 2 root-step
applications = (n-1) * m node x site updates) of one tree of that shape.

  value  device-resident: the program of the tree is already on the device, the step is the
         launch(es) of the Fitch kernel (+ the tiny count-finalise kernel), CUDA events on the
         engine's stream.
  e2e    through the C-ABI (lvb_gpu_tree_commit) with HOST topology arrays: host schedule
         compile, pinned H2D of the program, kernels, D2H of the tree length, every step, each
         step a different random tree.
  extra  SA trees scored/s: batches of candidate trees (NNI neighbourhood of the current tree)
         through lvb_gpu_score_batch, the other half of BASELINE.json's metric.

N > 1 (torchrun, one rank per GPU): BASELINE configs[4] style replicas — every rank holds the
alignment and evaluates its own trees, no data-path collective ("scaling": "weak").
`--mode siteblock` instead shards the words of ONE tree over the ranks with an NCCL all-reduce
of the partial lengths inside the C-ABI (BASELINE configs[3]); that is "strong" scaling.

--impl reference times the reference's own CPU getplen (oracle/_ref, OpenMP, all host cores)
on a bounded column sample of the same workload.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
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
    ap.add_argument("--mode", default="replicas", choices=["replicas", "siteblock"])
    ap.add_argument("--taxa", type=int, default=2000)
    ap.add_argument("--sites", type=int, default=1_000_000)
    ap.add_argument("--cpu-sample-sites", type=int, default=100_000)
    ap.add_argument("--candidates", type=int, default=512)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    return ap.parse_args()


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md)"


def traffic_from_profile():
    """DRAM bytes per launch of the dominant kernel from the committed `ncu --set full` capture
    (profiles/ncu_traffic.json, written by tools/ncu_summary.py); None if there is none."""
    p = os.path.join(ROOT, "profiles", "ncu_traffic.json")
    if not os.path.exists(p):
        return None, None
    d = json.load(open(p))
    return d.get("dram_bytes_per_pass"), d.get("source")


class ClockSampler:
    """SM clock and throttle reasons sampled DURING the timed regions (NVML, every 20 ms;
    nvidia-smi as a fallback)."""

    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index, self.samples, self.stop = index, [], threading.Event()
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
                try:
                    sm = nv.nvmlDeviceGetClockInfo(self.handle, nv.NVML_CLOCK_SM)
                    r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.handle)
                    self.samples.append([str(sm), str(self.max_sm)] + ["Active" if r & b else "Not Active" for _, b in bits])
                except Exception:
                    pass
                self.stop.wait(0.02)
            return
        while not self.stop.is_set():
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


def unpack_text(enc: np.ndarray, sites: int):
    """first `sites` sites of the packed one-hot alignment as text rows (for the CPU reference)"""
    lut = np.zeros(16, dtype=np.uint8)
    lut[[1, 2, 4, 8]] = np.frombuffer(b"ACGT", dtype=np.uint8)
    lut[15] = ord("N")
    words = (sites + 15) // 16
    rows = []
    for i in range(enc.shape[0]):
        b = enc[i, :words].view(np.uint8)
        nib = np.empty(b.size * 2, dtype=np.uint8)
        nib[0::2] = b & 15
        nib[1::2] = b >> 4
        rows.append(bytes(lut[nib[:sites]]))
    return rows


def cpu_reference_leg(enc, left, right, root, sample_sites, threads, seconds_budget=12.0):
    """Reference getplen (oracle/_ref) on a column sample: full passes, all host threads."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import oracles

    if not oracles.have_ref():
        return None
    n = enc.shape[0]
    rows = unpack_text(enc, sample_sites)
    msa = oracles.RefMSA.from_rows(rows, threads=threads)
    t = msa.tree()
    B = 2 * n - 3
    parent = np.full(B, -1, dtype=np.int64)
    for v in list(range(n, B)) + [root]:
        parent[left[v]] = v
        parent[right[v]] = v
    t.set_topology(parent, left, right, root)
    L = oracles.ref()
    import ctypes as C

    length = C.c_long(0)
    one = L.ref_time_full_passes(msa.h, t.t, root, 1, C.byref(length))  # warm-up + calibration
    reps = max(2, min(200, int(seconds_budget / max(one, 1e-4))))
    secs = L.ref_time_full_passes(msa.h, t.t, root, reps, C.byref(length))
    updates = (n - 1) * msa.m * reps
    return {
        "value": updates / secs, "unit": UNIT, "cores": int(msa.n_threads), "kind": "reference",
        "sample": f"{reps} full getplen passes (unmodified reference, OpenMP team of {msa.n_threads}) on the first "
                  f"{msa.m} of the sites of the same {n}-taxon alignment and tree; {secs / reps * 1e3:.2f} ms/pass",
        "ms_per_pass": secs / reps * 1e3, "length": int(length.value), "sites": int(msa.m),
    }


def nni_candidates(left, right, root, n, count, seed):
    """`count` NNI rearrangements of the tree with their dirty paths (host test-input code)"""
    from lvb_b200 import Topology

    B = 2 * n - 3
    parent = np.full(B, -1, dtype=np.int64)
    for v in list(range(n, B)) + [root]:
        parent[left[v]] = v
        parent[right[v]] = v
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
        out.append(Topology(l2, r2, root, dirty))
        index.append(2 * (u - n) + (1 if swap_a else 0))
    return out, index


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
        enc = synthetic.random_packed(n, (args.cpu_sample_sites + 15) // 16, seed=1)
        left, right, root = synthetic.random_topology(n, seed=3)
        t0 = time.time()
        # W warm-up + K timed "steps": each step is a bounded sample (one pass on the column sample)
        leg = cpu_reference_leg(enc, left, right, root, args.cpu_sample_sites, nproc,
                                seconds_budget=max(1.0, 0.05 * args.steps))
        if leg is None:
            print(json.dumps({"impl": "reference", "unavailable": "oracle/_ref not built on this box"}))
            return
        line = {
            "impl": "reference", "metric": METRIC, "value": leg["value"], "unit": UNIT, "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": leg["ms_per_pass"], "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "u64", "data": "synthetic",
            "config": {"workload": f"full getplen passes, {n} taxa x {leg['sites']} sites (column sample of the "
                                   f"{n} x {m} workload), reference CPU build, {leg['cores']} OpenMP threads"},
            "cpu_baseline": {k: leg[k] for k in ("value", "unit", "cores", "kind", "sample")},
            "e2e": {"value": leg["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "wall_s": time.time() - t0,
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

    from lvb_b200 import Engine, Topology

    siteblock = args.mode == "siteblock" and world > 1
    eng = Engine(local_rank)
    enc = synthetic.random_packed(n, nwords, seed=1)
    if siteblock:
        from lvb_b200 import sharding

        w0, wc = sharding.word_block(nwords, rank, world)
        eng.upload(enc, word_begin=w0, word_count=wc, max_trees=2)
        uid = [Engine.comm_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(uid, src=0)
        eng.comm_init(world, rank, uid[0])
    else:
        eng.upload(enc, max_trees=2)
    tree = eng.tree_create()
    # every rank its own trees in replica mode; the same tree on every rank in site-block mode
    tree_seed = 3 if siteblock else 3 + 1000 * rank
    left, right, root = synthetic.random_topology(n, seed=tree_seed)
    topo = Topology(left, right, root)
    updates_per_pass = (n - 1) * m  # real sites; the reference's count
    my_words = eng.word_count
    algo_bytes_per_pass = (3 * n - 6) * 8 * my_words  # 24 B per node-word update, SURVEY.md §8(d)

    length0 = eng.commit(tree, topo)
    warm = max(3, args.warmup)
    for _ in range(warm):
        eng.rerun_last_batch()
    eng.synchronize()

    with ClockSampler(local_rank) as clocks:
        # ---- region A: device-resident; CUDA events on the engine's own stream ----
        ext = torch.cuda.ExternalStream(eng.stream, device=torch.device("cuda", local_rank))
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        l0 = eng.launches
        eng.kernel_time(reset=True)
        barrier()
        ev0.record(ext)
        for _ in range(args.steps):
            eng.rerun_last_batch()
        ev1.record(ext)
        eng.synchronize()
        barrier()
        wall_a = max_over_ranks(ev0.elapsed_time(ev1) * 1e-3)
        k_ms, k_launch = eng.kernel_time()
        gpu_launches = eng.launches - l0
        kernel_ms = max_over_ranks(k_ms / max(1, k_launch))

        # ---- region B: end to end through the C-ABI, host topology arrays in, length out ----
        topos = [Topology(*synthetic.random_topology(n, seed=tree_seed + 17 * (s + 1))) for s in range(args.steps)]
        if siteblock:
            topos = [Topology(*synthetic.random_topology(n, seed=3 + 17 * (s + 1))) for s in range(args.steps)]
        for s in range(min(3, args.steps)):
            eng.commit(tree, topos[s])
        barrier()
        t0 = time.perf_counter()
        lens = [eng.commit(tree, tp) for tp in topos]
        barrier()
        wall_b = max_over_ranks(time.perf_counter() - t0)

        # ---- SA candidate scoring: trees scored/s ----
        eng.commit(tree, topo)
        cands, cand_index = nni_candidates(left, right, root, n, args.candidates, seed=5 + (0 if siteblock else rank))
        trees = [tree] * len(cands)
        got = eng.score_batch(trees, cands)
        barrier()
        t0 = time.perf_counter()
        reps_c = 3
        for _ in range(reps_c):
            got = eng.score_batch(trees, cands)
        barrier()
        wall_c = max_over_ranks(time.perf_counter() - t0)
        eng.kernel_time(reset=True)
        for _ in range(reps_c):
            eng.rerun_last_batch()
        eng.synchronize()
        c_ms, c_launch = eng.kernel_time()
        dirty_mean = float(np.mean([len(c.dirty) for c in cands]))
        # rows each candidate's program reads from HBM/L2 (operands >= 0), from the host compiler
        from lvb_b200 import compile_schedule

        rows_read = [int((compile_schedule(n, c, store=False)[0][:, :2] >= 0).sum()) for c in cands[:64]]
        rows_read_mean = float(np.mean(rows_read))

        # ---- exhaustive NNI neighbourhood (deterministic_hillclimb's 2(n-3) candidates) in one call ----
        scratch = eng.tree_create()
        nn = eng.nni_neighbourhood(tree, scratch, topo)
        barrier()
        t0 = time.perf_counter()
        reps_n = 3
        for _ in range(reps_n):
            nn = eng.nni_neighbourhood(tree, scratch, topo)
        barrier()
        wall_n = max_over_ranks(time.perf_counter() - t0)
        nn_kernel_ms = eng.last_kernel_ms()
        # the neighbourhood call and the dirty-path route must agree on the candidates they share
        nn_ok = bool(np.array_equal(nn[np.asarray(cand_index)], np.asarray(got)))

    total_ranks = 1 if siteblock else world
    value = total_ranks * args.steps * updates_per_pass / wall_a
    e2e_value = total_ranks * args.steps * updates_per_pass / wall_b
    peak, peak_src = peaks()
    achieved = algo_bytes_per_pass / (kernel_ms * 1e-3) / 1e9
    instr_bytes = (n - 1) * 16

    traffic, traffic_src = traffic_from_profile()
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": warm,
        "ms_per_step": wall_a / args.steps * 1e3, "higher_is_better": True,
        "scaling": "strong" if siteblock else "weak", "vs_baseline": None, "dtype": "u64 (4-bit state sets, 32-bit lanes)",
        "data": "synthetic",
        "config": {
            "workload": f"full getplen pass, {n} taxa x {m} sites (BASELINE.json configs[3] shape), "
                        + ("site blocks over ranks + NCCL all-reduce of lengths" if siteblock else
                           "one tree per step per GPU, alignment replicated per GPU"),
            "taxa": n, "sites": m, "words_per_row": nwords, "state_bytes_per_tree": (2 * n - 3) * 8 * nwords,
            "l2": "inputs larger than L2: 1.0 GB alignment + 2.0 GB state rows per pass vs 126 MB L2",
            "parallelism": f"{args.mode}{world}",
        },
        "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": instr_bytes + 64, "d2h_bytes_per_step": 8,
                "ms_per_step": wall_b / args.steps * 1e3,
                "api": "lvb_gpu_tree_commit(host left/right/root -> length), a different random tree every step"},
        "gpu_launches": gpu_launches,
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                     "traffic": traffic, "traffic_source": traffic_src, "traffic_unit": "DRAM bytes per pass (ncu)",
                     "traffic_gbs": (traffic / (kernel_ms * 1e-3) / 1e9) if traffic else None,
                     "traffic_frac": (traffic / (kernel_ms * 1e-3) / 1e9 / peak) if traffic else None,
                     "peak_source": peak_src, "kernel": "fitch_exec_kernel<true> (one launch per pass)",
                     "timing": "CUDA events on the engine's stream around each launch, averaged over the timed steps",
                     "kernel_ms_per_pass": kernel_ms,
                     "algorithmic_bytes_per_pass": algo_bytes_per_pass,
                     "note": "algorithmic = 24 B per node-word update (2 reads + 1 write); the kernel keeps "
                             "intermediate rows on chip, so its real DRAM traffic is ~16 B per node-word"},
        "clocks": clocks.summary(),
        "sa_trees_scored_per_s": {
            "e2e": total_ranks * reps_c * len(cands) / wall_c,
            "device": total_ranks * len(cands) / (c_ms / max(1, c_launch) * 1e-3),
            "batch": len(cands), "mean_dirty_nodes": dirty_mean,
            "api": "lvb_gpu_score_batch: NNI candidates of the current tree, one launch per batch",
        },
        "roofline_score": {
            "bound": "hbm (L2-assisted)", "unit": "GB/s", "peak": peak,
            "achieved": rows_read_mean * 8 * my_words * len(cands) / (c_ms / max(1, c_launch) * 1e-3) / 1e9,
            "frac": rows_read_mean * 8 * my_words * len(cands) / (c_ms / max(1, c_launch) * 1e-3) / 1e9 / peak,
            "rows_read_per_candidate": rows_read_mean,
            "kernel": "fitch_score_kernel", "kernel_ms_per_batch": c_ms / max(1, c_launch),
            "note": "SURVEY.md 8(d): 8 B x nwords x (clean rows read) per candidate, nothing written; candidates "
                    "of one tree share sibling rows, so part of this is served by L2 and the figure may exceed the "
                    "HBM peak",
        },
        "nni_neighbourhood": {
            "candidates": int(len(nn)), "ms_per_call": wall_n / reps_n * 1e3,
            "candidates_per_s": total_ranks * len(nn) / (wall_n / reps_n),
            "four_row_kernel_ms": nn_kernel_ms, "agrees_with_score_batch": bool(nn_ok),
            "api": "lvb_gpu_nni_neighbourhood: up-pass program + four-row kernel, host arrays in, 2(n-3) lengths out",
        },
        "length_check": int(length0), "e2e_lengths_head": [int(x) for x in lens[:3]],
    }
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        try:
            cw = (args.cpu_sample_sites + 15) // 16
            leg = cpu_reference_leg(enc[:, :cw].copy(), left, right, root, args.cpu_sample_sites, nproc)
            if leg is not None:
                line["cpu_baseline"] = {k: leg[k] for k in ("value", "unit", "cores", "kind", "sample")}
        except Exception as exc:  # the baseline is reported, never required
            line["cpu_baseline"] = {"value": None, "unit": UNIT, "cores": nproc, "kind": "reference",
                                    "sample": f"failed: {exc}"}
    if rank == 0:
        print(json.dumps(line))
    eng.close()
    if dist is not None:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
