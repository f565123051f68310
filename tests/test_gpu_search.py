"""The search on the real device: `lvb_gpu` (C++ host layer -> C-ABI -> CUDA kernels) against runs
of the unmodified reference — recorded in tests/golden, and, where oracle/_ref travelled to this
box, produced here by the reference binary itself.  Parity bar: identical integers and identical
files (every accepted/rejected move of changeAccepted.tsv, the Newick output)."""
import json
import os
import subprocess
import sys
import tempfile

import numpy as np
import pytest

import cli_harness as H
import oracles

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))
CLI_RUNS = json.load(open(os.path.join(HERE, "golden", "cli_runs.json")))
SA_RUNS = json.load(open(os.path.join(HERE, "golden", "sa_runs.json")))

GPU_CASES = ["test_treelength_1@1", "test_treelength_2_a2@1", "test_treelength_4@2", "test_treelength_5_thread_2",
             "test_treelength_6_thread_2", "test_treelength_6_thread_3", "test_trees_recovered_4@1"]


@pytest.mark.parametrize("case", GPU_CASES)
def test_reference_executable_tests_on_gpu(case):
    want = CLI_RUNS[case]
    got = H.run_cli(H.cli_input(case.split("@")[0]), want["flags"], mock=False)
    H.check_against(got, want, case)
    assert got["kernel_launches"] > 0


@pytest.mark.parametrize("algo", [0, 1, 2])
def test_example_trajectory_on_gpu(algo):
    """BASELINE.json configs[0]"""
    want = SA_RUNS[f"a{algo}"]
    got = H.run_cli(H.example_input(), ["-s", "12345", "-a", str(algo)], mock=False)
    H.check_against(got, want, f"example a={algo}")


def test_config2_trajectory_on_gpu():
    """BASELINE.json configs[1]: synthetic 100 taxa x 10k sites, deterministic SA trajectory vs reference"""
    want = SA_RUNS["config2_a1"]
    got = H.run_cli(H.config2_input(), want["flags"], mock=False)
    H.check_against(got, want, "config 2")
    assert got["score"] == 96403 and got["rearrangements"] == 241917


@pytest.mark.skipif(not os.path.exists(oracles.REF_BIN), reason="oracle/_ref/lvb_ref not on this box")
def test_same_files_as_the_reference_binary_run_here():
    """reference and lvb_gpu side by side on this box (same libm): byte-identical outputs"""
    raw = H.example_input()
    with tempfile.TemporaryDirectory() as td:
        open(os.path.join(td, "infile"), "wb").write(raw)
        p = subprocess.run([oracles.REF_BIN, "-s", "777", "-a", "1", "-p", "1", "-v"], cwd=td, capture_output=True,
                           text=True, timeout=600)
        assert p.returncode == 0
        ref_tsv, ref_out = H.md5_file(os.path.join(td, "changeAccepted.tsv")), H.md5_file(os.path.join(td, "outtree"))
    got = H.run_cli(raw, ["-s", "777", "-a", "1"], mock=False)
    assert got["rc"] == 0, got["stdout"][-500:]
    assert (got["tsv_md5"], got["outtree_md5"]) == (ref_tsv, ref_out)


def test_replicates_on_gpu():
    """BASELINE.json configs[4] in small: 8 chains in lock-step batches == 8 single-chain runs"""
    def run(seed, reps):
        p = subprocess.run([sys.executable, os.path.join(HERE, "run_host_search.py"), "example.npz:text", str(seed), "1",
                            "0", str(reps), "30000", "4"], capture_output=True, text=True, timeout=900)
        assert p.returncode == 0, p.stderr[-2000:]
        return json.loads(p.stdout.strip().splitlines()[-1])

    multi = run(100, 8)
    for r in (0, 3, 7):
        single = run(100 + r, 1)
        for k in ("final_length", "rearrangements", "topologies", "initial_length", "start_temperature"):
            assert multi["results"][r][k] == single["results"][0][k], (r, k)
        assert multi["outtrees"][r] == single["outtrees"][0]
    # proposals of several chains (and several proposals per chain) share a launch
    assert multi["results"][0]["score_launches"] < sum(r["trees_scored"] for r in multi["results"])


def test_compatible_getplen_on_gpu(golden_dir):
    """getplen() with the reference's signature and dirty-flag protocol (one tree per call):
    treecopy + mutate + getplen, accept by pointer swap, as Anneal does (Solve.c:288-323)"""
    from lvb_b200 import host

    g = np.load(os.path.join(golden_dir, "example.npz"))
    enc = g["enc"]
    h = host.HostMSA.from_rows([bytes(r) for r in g["text"]])
    L = host.load_host_library()
    assert L.lvbh_device_open(h.h, 0, 4) == 0, L.lvbh_last_error()
    try:
        h.rinit(12345)
        cur = h.tree().random()
        assert cur.getplen() == 11808
        assert cur.getplen() == 11808  # nothing dirty: cached counts + root step
        prop = h.tree()
        want = {"nni": 11834, "spr": 11835, "tbr": 11827}
        for kind in ("nni", "spr", "tbr"):
            prop.mutate(kind, cur)
            assert prop.getplen() == want[kind]
            cur, prop = prop, cur
        cur.arbreroot()
        assert cur.getplen() == 11827
        # a longer run against the oracle, with rejections (the proposal buffer is reused)
        for i in range(60):
            prop.mutate(("nni", "spr", "tbr")[i % 3], cur)
            _, left, right, _ = prop.get()
            length, _, _ = oracles.oracle_getplen(enc, left, right, prop.root)
            assert prop.getplen() == length, i
            if i % 3 == 0:
                cur, prop = prop, cur
    finally:
        L.lvbh_device_close()


def test_matchange_on_the_device(monkeypatch):
    """the example alignment loses 16 constant columns (1000 -> 984): with
    LVB_GPU_MATCHANGE=1 the scan and MinimumTreeLength come from lvb_gpu_column_stats and the
    whole run (starting temperature uses min_len_tree) is unchanged"""
    monkeypatch.setenv("LVB_GPU_MATCHANGE", "1")
    want = SA_RUNS["a1"]
    got = H.run_cli(H.example_input(), ["-s", "12345", "-a", "1"], mock=False)
    H.check_against(got, want, "example a=1, matchange on the device")
    assert "In total, 16 columns are excluded" in got["stdout"]


@pytest.mark.parametrize("env", [{"LVB_GPU_INLINE": "0"}, {"LVB_GPU_FUSE": "0", "LVB_GPU_DEEP": "0"},
                                 {"LVB_GPU_INLINE": "0", "LVB_GPU_FUSE": "0", "LVB_GPU_DEEP": "0"}])
def test_staged_and_unfused_kernel_routes(monkeypatch, env):
    """small batches normally ride in the kernel parameters, finalize inside the Fitch kernel and use
    the deep pipeline; with the switches off the same search goes through the staged copy, the
    separate finalize launch and the 3-deep / 1-deep kernels — same trajectory either way"""
    for k, v in env.items():
        monkeypatch.setenv(k, v)
    want = CLI_RUNS["test_treelength_6_thread_2"]
    got = H.run_cli(H.cli_input("test_treelength_6_thread_2"), want["flags"], mock=False)
    H.check_against(got, want, str(env))


def test_driver_groups_on_gpu():
    """8 chains in 2 driver groups (second context = device-to-device clone of the leaf rows,
    lvb_gpu_matrix_clone; programs compiled in the chains' loop, lvb_gpu_*_pre) == single-chain runs"""
    def run(seed, reps, groups):
        p = subprocess.run([sys.executable, os.path.join(HERE, "run_host_search.py"), "treelength.npz:tl6_thread_2_text",
                            str(seed), "1", "0", str(reps), "0", "4", str(groups)], capture_output=True, text=True, timeout=900)
        assert p.returncode == 0, p.stderr[-2000:]
        return json.loads(p.stdout.strip().splitlines()[-1])

    multi = run(40, 8, 2)
    assert multi["results"][0]["groups"] == 2
    for r in (0, 5, 7):
        single = run(40 + r, 1, 0)
        for k in ("final_length", "rearrangements", "topologies", "initial_length", "start_temperature"):
            assert multi["results"][r][k] == single["results"][0][k], (r, k)
        assert multi["outtrees"][r] == single["outtrees"][0]
