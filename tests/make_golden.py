"""Generate tests/golden/*.npz|json from the UNMODIFIED reference.

Run in the build container only (it reads /root/reference and oracle/_ref, built by
`make -C oracle ref`):    python tests/make_golden.py
The GPU box has no /root/reference; tests there read the committed fixtures.

What is recorded (all produced by reference code through oracle/ref_shim.cpp or the
reference CLI binary oracle/_ref/lvb_ref):
  example.npz      examples/inputs 100x1000 alignment after matchange (984 columns): text rows,
                   DNAToBinary words, MinimumTreeLength; the SURVEY §8(c) known-answer chain
                   (rinit 12345 -> PullRandomTree -> getplen -> nni -> spr -> tbr -> arbreroot)
                   with topology, dirty flags, length, per-node changes and all state rows.
  proposals.npz    300 proposals from an evolving current tree on the example alignment
                   (nni/spr/tbr in turn, accepted when not worse, re-rooted every 50): the
                   candidate topology, its dirty set and the reference getplen length.
  treelength.npz   the alignments of test/src/COMMON/test_treelength_* with the score each
                   runtest pins, a random tree and 40 proposals each.
  rng.npz          uni()/randpint() streams for several seeds.
  sa_runs.json     end-to-end runs of the reference CLI on the example (seed 12345, -a 0/1/2).
"""
from __future__ import annotations

import hashlib
import json
import os
import re
import subprocess
import sys
import tempfile

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import oracles  # noqa: E402

REF = "/root/reference"
OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
EXAMPLE = f"{REF}/examples/inputs/simulated_matrix_100x1000_phylip.phy"


def snap(tree: oracles.RefTree):
    parent, left, right, changes, dirty = tree.get()
    return dict(parent=parent, left=left, right=right, changes=changes, dirty=dirty, root=tree.root)


def text_matrix(msa: oracles.RefMSA) -> np.ndarray:
    return np.frombuffer(b"".join(msa.rows_text()), dtype=np.uint8).reshape(msa.n, msa.m).copy()


def gen_example():
    L = oracles.ref()
    msa = oracles.RefMSA.load(EXAMPLE)
    out = dict(text=text_matrix(msa), enc=msa.enc(), min_len=msa.min_len, original_m=msa.original_m)
    L.ref_rinit(12345)
    cur = msa.tree().random()
    steps = []
    s = snap(cur)
    s["length"] = cur.getplen()
    s["length_again"] = cur.getplen()
    after = snap(cur)
    s["changes_after"] = after["changes"]
    s["rows_after"] = cur.rows()
    steps.append(("random", s))
    prev = cur
    for kind in ("nni", "spr", "tbr"):
        nxt = msa.tree().mutate(kind, prev)
        s = snap(nxt)
        s["length"] = nxt.getplen()
        s["changes_after"] = snap(nxt)["changes"]
        s["rows_after"] = nxt.rows()
        steps.append((kind, s))
        prev = nxt
    prev.arbreroot()
    s = snap(prev)
    s["length"] = prev.getplen()
    s["changes_after"] = snap(prev)["changes"]
    s["rows_after"] = prev.rows()
    steps.append(("reroot", s))
    out["next_uni"] = L.ref_uni()
    for name, s in steps:
        for k, v in s.items():
            out[f"{name}_{k}"] = v
    np.savez_compressed(os.path.join(OUT, "example.npz"), **out)
    print("example:", [int(s["length"]) for _, s in steps], "next uni", out["next_uni"])


def proposal_series(msa: oracles.RefMSA, seed: int, count: int, reroot_every: int):
    L = oracles.ref()
    L.ref_rinit(seed)
    cur = msa.tree().random()
    cur_len = cur.getplen()
    prop = msa.tree()
    recs = dict(kind=[], left=[], right=[], root=[], dirty=[], length=[], accepted=[], cur_len=[])
    start = snap(cur)
    kinds = ("nni", "spr", "tbr")
    for i in range(count):
        if reroot_every and i and i % reroot_every == 0:
            cur.arbreroot()  # leaves every internal node of `cur` dirty, like Solve.c:240-247
        k = kinds[i % 3]
        prop.mutate(k, cur)
        s = snap(prop)
        length = prop.getplen()
        recs["kind"].append(i % 3)
        recs["left"].append(s["left"])
        recs["right"].append(s["right"])
        recs["root"].append(prop.root)
        recs["dirty"].append(s["dirty"])
        recs["length"].append(length)
        recs["cur_len"].append(cur_len)
        acc = length <= cur_len
        recs["accepted"].append(int(acc))
        if acc:
            cur, prop = prop, cur
            cur_len = length
    out = {k: np.array(v) for k, v in recs.items()}
    out["start_left"], out["start_right"], out["start_root"] = start["left"], start["right"], start["root"]
    out["left"] = out["left"].astype(np.int32)
    out["right"] = out["right"].astype(np.int32)
    out["dirty"] = out["dirty"].astype(np.uint8)
    return out


def gen_proposals():
    msa = oracles.RefMSA.load(EXAMPLE)
    out = proposal_series(msa, 4242, 300, 50)
    out["seed"] = 4242
    np.savez_compressed(os.path.join(OUT, "proposals.npz"), **out)
    print("proposals: accepted", int(out["accepted"].sum()), "of", len(out["accepted"]),
          "full passes", int((out["dirty"][:, msa.n:].sum(axis=1) == msa.n - 3).sum()))


TREELENGTH = {  # directory -> (score pinned by its runtest, CLI flags used there)
    "test_treelength_1": (1, ""),
    "test_treelength_2": (5, ""),
    "test_treelength_3": (4, ""),
    "test_treelength_4": (9, ""),
    "test_treelength_5_thread_2": (1846, "-s 509739986 -p 2"),
    "test_treelength_6_thread_2": (1628, "-p 2 -s 509739986"),
    "test_treelength_6_thread_3": (297, "-p 3 -s 465380177"),
    "test_treelength_7_thread_2": (1006, "-p 2"),
}


def gen_treelength():
    out = {}
    for d, (score, flags) in TREELENGTH.items():
        path = f"{REF}/test/src/COMMON/{d}/infile"
        runtest = open(f"{REF}/test/src/COMMON/{d}/runtest").read()
        assert re.search(rf"Tree score: \+{score}\\n", runtest), (d, score)
        msa = oracles.RefMSA.load(path)
        key = d.replace("test_treelength_", "tl")
        out[f"{key}_text"] = text_matrix(msa)
        out[f"{key}_enc"] = msa.enc()
        out[f"{key}_min_len"] = msa.min_len
        out[f"{key}_original_m"] = msa.original_m
        out[f"{key}_score"] = score
        ser = proposal_series(msa, 7, 40, 10)
        for k, v in ser.items():
            out[f"{key}_{k}"] = v
        print(d, "n,m =", msa.n, msa.m, "nwords", msa.nwords, "pinned score", score)
    np.savez_compressed(os.path.join(OUT, "treelength.npz"), **out)


def gen_rng():
    L = oracles.ref()
    out = {}
    for seed in (0, 1, 12345, 509739986, 900000000):
        L.ref_rinit(seed)
        out[f"uni_{seed}"] = np.array([L.ref_uni() for _ in range(2000)])
        L.ref_rinit(seed)
        ups = [1, 2, 3, 7, 96, 193, 1000, 123456789]
        out[f"randpint_{seed}"] = np.array([[L.ref_randpint(u) for u in ups] for _ in range(200)])
        out["randpint_uppers"] = np.array(ups)
    np.savez_compressed(os.path.join(OUT, "rng.npz"), **out)


def gen_sa_runs():
    runs = {}
    for a in (0, 1, 2):
        with tempfile.TemporaryDirectory() as td:
            p = subprocess.run([oracles.REF_BIN, "-i", EXAMPLE, "-s", "12345", "-a", str(a), "-p", "1", "-v"],
                               cwd=td, capture_output=True, text=True, check=True)
            o = p.stdout
            tsv = open(os.path.join(td, "changeAccepted.tsv")).read().splitlines()
            runs[f"a{a}"] = dict(
                score=int(re.search(r"Tree score: +(\d+)", o).group(1)),
                topologies=int(re.search(r"Topologies recovered: +(\d+)", o).group(1)),
                rearrangements=int(re.search(r"Rearrangements evaluated: +(\d+)", o).group(1)),
                start_temp=re.search(r"Starting Temperature: +([0-9.]+)", o).group(1),
                outtree_md5=hashlib.md5(open(os.path.join(td, "outtree"), "rb").read()).hexdigest(),
                tsv_md5=hashlib.md5(("\n".join(tsv) + "\n").encode()).hexdigest(),
                tsv_head=tsv[:41],
                init_len=int(open(os.path.join(td, "sum")).read().splitlines()[1].split("\t")[2]),
            )
            print("sa run a=%d:" % a, {k: v for k, v in runs[f"a{a}"].items() if k != "tsv_head"})
    json.dump(runs, open(os.path.join(OUT, "sa_runs.json"), "w"), indent=1)


def run_reference_cli(infile_bytes: bytes, flags, fmt_flags=()):
    """the reference binary in a directory of its own, verbose (-v writes changeAccepted.tsv and sum)"""
    with tempfile.TemporaryDirectory() as td:
        open(os.path.join(td, "infile"), "wb").write(infile_bytes)
        p = subprocess.run([oracles.REF_BIN, *flags, "-v"], cwd=td, capture_output=True, text=True, check=True)
        o = p.stdout
        tsv = open(os.path.join(td, "changeAccepted.tsv")).read().splitlines()
        return dict(
            score=int(re.search(r"Tree score: +(\d+)", o).group(1)),
            topologies=int(re.search(r"Topologies recovered: +(\d+)", o).group(1)),
            rearrangements=int(re.search(r"Rearrangements evaluated: +(\d+)", o).group(1)),
            start_temp=re.search(r"Starting Temperature: +([0-9.]+)", o).group(1),
            outtree_md5=hashlib.md5(open(os.path.join(td, "outtree"), "rb").read()).hexdigest(),
            outtree=open(os.path.join(td, "outtree")).read() if os.path.getsize(os.path.join(td, "outtree")) < 20000 else None,
            tsv_md5=hashlib.md5(("\n".join(tsv) + "\n").encode()).hexdigest(),
            tsv_rows=len(tsv) - 1,
            init_len=int(open(os.path.join(td, "sum")).read().splitlines()[1].split("\t")[2]),
            flags=list(flags),
        )


# the reference's executable tests that pin the path: directory -> (flags of its runtest, pinned regexes)
CLI_CASES = {
    "test_treelength_1": ([], {"score": 1}),
    "test_treelength_2": ([], {"score": 5}),
    "test_treelength_2_a1": (["-a1"], {"score": 5}),
    "test_treelength_2_a2": (["-a2"], {"score": 5}),
    "test_treelength_3": ([], {"score": 4}),
    "test_treelength_4": ([], {"score": 9}),
    "test_treelength_5_thread_2": (["-s", "509739986", "-p", "2"], {"score": 1846}),
    "test_treelength_6_thread_2": (["-p", "2", "-s", "509739986"], {"score": 1628}),
    "test_treelength_6_thread_3": (["-p", "3", "-s", "465380177"], {"score": 297}),
    "test_treelength_7_thread_2": (["-p", "2"], {"score": 1006}),
    "test_trees_recovered_1": ([], {"topologies": 15}),
    "test_trees_recovered_2": ([], {"topologies": 3}),
    "test_trees_recovered_3": ([], {"topologies": 1}),
    "test_trees_recovered_4": ([], {"topologies": 105}),
    "test_trees_recovered_6": ([], {"topologies": 1}),
    "test_trees_recovered_7": ([], {"topologies": 1}),
}


def gen_cli_cases():
    """Inputs of the reference's own executable tests + what the unmodified reference prints and
    writes for them; seed-free tests are recorded at seeds 1 and 2 (their pins hold for any seed)."""
    inputs, runs = {}, {}
    for d, (flags, pins) in CLI_CASES.items():
        raw = open(f"{REF}/test/src/COMMON/{d}/infile", "rb").read()
        inputs[d] = np.frombuffer(raw, dtype=np.uint8)
        seeded = "-s" in flags
        for seed in ([None] if seeded else [1, 2]):
            fl = list(flags) + ([] if seeded else ["-s", str(seed)])
            r = run_reference_cli(raw, fl)
            for k, v in pins.items():
                assert r[k] == v, (d, k, r[k], v)
            r["pins"] = pins
            runs[d + ("" if seeded else f"@{seed}")] = r
            print(d, fl, {k: r[k] for k in ("score", "topologies", "rearrangements", "start_temp")})
    np.savez_compressed(os.path.join(OUT, "cli_inputs.npz"), **inputs)
    json.dump(runs, open(os.path.join(OUT, "cli_runs.json"), "w"), indent=1)


def gen_config2():
    """BASELINE.json configs[1]: synthetic 100 x 10k (SURVEY.md appendix A generator), reference run"""
    sys.path.insert(0, os.path.dirname(OUT.rstrip("/")).rsplit("/tests", 1)[0])
    from lvb_b200 import synthetic

    with tempfile.TemporaryDirectory() as td:
        path = os.path.join(td, "c2.phy")
        synthetic.write_phylip(path, synthetic.survey_alignment(100, 10000, 1, 0.05))
        raw = open(path, "rb").read()
    assert hashlib.md5(raw).hexdigest() == "a11fa9393aaf6c6b8150b4491e8a4d90"
    r = run_reference_cli(raw, ["-s", "12345", "-a", "1", "-p", "1"])
    r["input_md5"] = hashlib.md5(raw).hexdigest()
    assert r["score"] == 96403 and r["rearrangements"] == 241917, r
    runs = json.load(open(os.path.join(OUT, "sa_runs.json")))
    runs["config2_a1"] = r
    json.dump(runs, open(os.path.join(OUT, "sa_runs.json"), "w"), indent=1)
    print("config 2:", {k: v for k, v in r.items() if k not in ("outtree",)})


def gen_example_raw():
    raw = open(EXAMPLE, "rb").read()
    np.savez_compressed(os.path.join(OUT, "example_phy.npz"), raw=np.frombuffer(raw, dtype=np.uint8))


if __name__ == "__main__":
    os.makedirs(OUT, exist_ok=True)
    if len(sys.argv) > 1:
        for fn in sys.argv[1:]:
            globals()[fn]()
        sys.exit(0)
    assert oracles.have_ref(), "build the reference first: make -C oracle ref"
    gen_example()
    gen_proposals()
    gen_treelength()
    gen_rng()
    gen_sa_runs()
    gen_example_raw()
    gen_cli_cases()
    gen_config2()
