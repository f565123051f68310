#!/usr/bin/env python
"""SA end to end on the GPU box: `lvb_gpu` next to the unmodified reference (`oracle/_ref`).

  configs 1, 2   complete searches, wall clock, same seed: lvb_gpu vs lvb_ref -p 1 / -p nproc;
                 score, rearrangements and outtree must be identical.
  configs 3, 5   a fixed budget of rearrangements on the GPU (-x), the reference's proposal loop
                 (mutate_* + getplen, Solve.c:288-300) timed on a bounded number of proposals
                 through oracle/_ref's shim.
Writes one JSON document (default gpurun_out/sa_bench.json).  Test/bench infrastructure: the
oracle is only the baseline and the checker here.
"""
from __future__ import annotations

import ctypes as C
import hashlib
import json
import os
import re
import subprocess
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
CLI = os.path.join(ROOT, "lvb_b200", "bin", "lvb_gpu")

from lvb_b200 import synthetic  # noqa: E402
import cli_harness as H  # noqa: E402
import oracles  # noqa: E402


def parse(o: str) -> dict:
    def num(pat, cast=int):
        m = re.search(pat, o)
        return cast(m.group(1)) if m else None

    return dict(score=num(r"Tree score: +(\d+)\n"), topologies=num(r"Topologies recovered: +(\d+)\n"),
                rearrangements=num(r"Rearrangements evaluated: +(\d+)\n"),
                trees_scored=num(r"trees scored: (\d+)"), score_launches=num(r"trees scored: \d+ in (\d+) launches"),
                kernel_launches=num(r"kernel launches: (\d+)"),
                search_seconds=num(r"Wall time of the search: +([0-9.]+) s", float),
                host_split=num(r"Host time split \(s\): +(.*)\n", str))


def run(binary: str, td: str, flags) -> dict:
    t0 = time.time()
    p = subprocess.run([binary, *flags], cwd=td, capture_output=True, text=True, timeout=3000)
    wall = time.time() - t0
    if p.returncode != 0:
        return dict(rc=p.returncode, tail=(p.stdout + p.stderr)[-600:])
    out = parse(p.stdout)
    out.update(rc=0, wall_s=wall, outtree_md5=H.md5_file(os.path.join(td, "outtree")))
    return out


def complete_search(name: str, raw: bytes, flags, nproc: int, with_ref: bool = True) -> dict:
    res = dict(config=name, flags=" ".join(flags))
    with tempfile.TemporaryDirectory() as td:
        open(os.path.join(td, "infile"), "wb").write(raw)
        res["lvb_gpu"] = run(CLI, td, flags)
        if with_ref and os.path.exists(oracles.REF_BIN):
            for p in sorted({1, nproc}):
                r = run(oracles.REF_BIN, td, [*flags, "-p", str(p)])
                res[f"reference_p{p}"] = r
                g = res["lvb_gpu"]
                r["identical_to_lvb_gpu"] = all(r.get(k) == g.get(k) for k in ("score", "rearrangements", "topologies", "outtree_md5"))
    g = res["lvb_gpu"]
    if g.get("rc") == 0 and g.get("search_seconds"):
        res["gpu_trees_scored_per_s"] = g["trees_scored"] / g["search_seconds"]
        # proposals of the reference's trajectory (what both programs have to get through), per second
        res["gpu_rearrangements_per_s"] = g["rearrangements"] / g["search_seconds"]
    return res


def budget_search(name: str, n: int, m: int, p: float, chains: int, budget: int, ref_proposals: int, nproc: int) -> dict:
    """GPU: -R chains, -x budget rearrangements per chain.  Reference: its proposal loop on the same alignment."""
    res = dict(config=name, taxa=n, sites=m, chains=chains, rearrangements_per_chain=budget)
    st = synthetic.states(n, m, 1, p)
    with tempfile.TemporaryDirectory() as td:
        synthetic.write_phylip(os.path.join(td, "infile"), st)
        g = run(CLI, td, ["-s", "1000", "-a", "1", "-R", str(chains), "-x", str(budget)])
        res["lvb_gpu"] = g
        if g.get("rc") == 0 and g.get("search_seconds"):
            # trees_scored is chain 0's count; every chain does the same number of evaluations +-
            res["gpu_trees_scored_per_s_all_chains"] = chains * g["trees_scored"] / g["search_seconds"]
            # speculative proposals that were scored but not used are in trees_scored, not here
            res["gpu_rearrangements_per_s_all_chains"] = chains * budget / g["search_seconds"]
    if oracles.have_ref():
        rows = synthetic.text_rows(st)
        L = oracles.ref()
        for threads in sorted({1, nproc}):
            msa = oracles.RefMSA.from_rows(rows, threads=threads)
            L.ref_rinit(12345)
            cur = msa.tree().random()
            cur.getplen()
            prop = msa.tree()
            gp, dirty = C.c_double(0), C.c_long(0)
            secs = L.ref_time_proposals(msa.h, cur.t, prop.t, cur.root, 3, ref_proposals, None, C.byref(gp), C.byref(dirty))
            res[f"reference_p{threads}"] = dict(
                proposals=ref_proposals, seconds=secs, trees_scored_per_s=ref_proposals / secs,
                getplen_share=gp.value / secs, mean_dirty=dirty.value / ref_proposals, threads=int(msa.n_threads),
                what="mutate_nni/mutate_spr alternating + getplen from one evaluated random tree (Solve.c:288-300)")
            del prop, cur, msa
    return res


def hillclimb(name: str, n: int, m: int, p: float, seed: int, with_ref: bool) -> dict:
    """deterministic_hillclimb (Solve.c:64-142) from the same random start tree: the host loop over
    lvb_gpu_nni_neighbourhood against the reference's own function (oracle/_ref), each in a process of its own"""
    res = dict(config=name, taxa=n, sites=m)
    runner = os.path.join(ROOT, "tests", "run_hillclimb.py")
    for key, flag in (("lvb_gpu", None), ("reference", "--ref")):
        if flag and not (with_ref and oracles.have_ref()):
            continue
        cmd = [sys.executable, runner] + ([flag] if flag else []) + [f"synthetic:{n}:{m}:{p}", str(seed)]
        t0 = time.time()
        q = subprocess.run(cmd, capture_output=True, text=True, timeout=3000)
        if q.returncode != 0:
            res[key] = dict(rc=q.returncode, tail=(q.stdout + q.stderr)[-400:])
            continue
        res[key] = json.loads(q.stdout.strip().splitlines()[-1])
        res[key]["wall_s"] = time.time() - t0
    if "reference" in res and "seconds" in res.get("lvb_gpu", {}) and "seconds" in res["reference"]:
        g, r = res["lvb_gpu"], res["reference"]
        res["identical"] = all(g[k] == r[k] for k in ("length", "iterations", "trees", "outtree_md5"))
        res["speedup"] = r["seconds"] / g["seconds"]
    return res


def main() -> None:
    out_path = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "gpurun_out", "sa_bench.json")
    which = sorted(set(sys.argv[2].split(","))) if len(sys.argv) > 2 else ["1", "2", "3", "5"]
    nproc = os.cpu_count() or 1
    os.makedirs(os.path.dirname(os.path.abspath(out_path)), exist_ok=True)
    if len(which) > 1:
        # one process per config: the reference keeps file-static buffers sized by the first
        # alignment it sees (TreeOperations.c), so two shapes must not share a process
        doc = dict(host_cores=nproc, results=[])
        for w in which:
            part = out_path + f".part{w}"
            p = subprocess.run([sys.executable, os.path.abspath(__file__), part, w])
            if p.returncode == 0 and os.path.exists(part):
                doc["results"] += json.load(open(part))["results"]
                os.remove(part)
            else:
                doc["results"].append(dict(config=f"config{w}", failed=p.returncode))
        json.dump(doc, open(out_path, "w"), indent=1)
        return
    doc = dict(host_cores=nproc, results=[])
    if "1" in which:
        for a in ("0", "1", "2"):
            doc["results"].append(complete_search(f"config1 example 100x984 a={a}", H.example_input(), ["-s", "12345", "-a", a], nproc))
    if "2" in which:
        doc["results"].append(complete_search("config2 synthetic 100x10000 a=1", H.config2_input(), ["-s", "12345", "-a", "1"], nproc))
    if "3" in which:
        doc["results"].append(budget_search("config3 synthetic 500x100000, 1 chain", 500, 100000, 0.02, 1, 20000, 100, nproc))
    if "5" in which:
        doc["results"].append(budget_search("config5 synthetic 1000x200000, 64 chains on one GPU", 1000, 200000, 0.02, 64, 3000, 40, nproc))
    if "h" in which:
        doc["results"].append(hillclimb("hillclimb synthetic 200x20000", 200, 20000, 0.05, 3, True))
        doc["results"].append(hillclimb("hillclimb synthetic 500x100000 (config 3 shape)", 500, 100000, 0.02, 3, False))
    json.dump(doc, open(out_path, "w"), indent=1)
    for r in doc["results"]:
        print(json.dumps(r))


if __name__ == "__main__":
    main()
