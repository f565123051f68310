"""The search (GetSoln / StartingTemperature / Anneal re-driven in speculative batches,
lvb_b200/host/solve.cpp) against runs of the UNMODIFIED reference recorded in tests/golden:
final score, number of topologies, number of rearrangements, starting temperature, initial
length, every line of changeAccepted.tsv (md5) and the Newick output (md5).

No GPU here: the `lvb_gpu` binary runs with tests/mock_gpu.c preloaded, so lengths come from
the CPU oracle and what is under test is the host logic.  test_gpu_search.py repeats the checks
with the real library on a B200."""
import json
import os
import subprocess
import sys

import pytest

import cli_harness as H

HERE = os.path.dirname(os.path.abspath(__file__))
CLI_RUNS = json.load(open(os.path.join(HERE, "golden", "cli_runs.json")))
SA_RUNS = json.load(open(os.path.join(HERE, "golden", "sa_runs.json")))

# the reference's own executable tests (test/src/COMMON), a selection that keeps the CPU suite short
CPU_CASES = ["test_treelength_1@1", "test_treelength_2_a1@2", "test_treelength_2_a2@1", "test_treelength_3@1",
             "test_treelength_4@2", "test_treelength_5_thread_2", "test_treelength_6_thread_3",
             "test_trees_recovered_1@2", "test_trees_recovered_4@1", "test_trees_recovered_7@2"]


@pytest.mark.parametrize("case", CPU_CASES)
def test_reference_executable_tests(case):
    want = CLI_RUNS[case]
    got = H.run_cli(H.cli_input(case.split("@")[0]), want["flags"], mock=True)
    H.check_against(got, want, case)
    for k, v in want["pins"].items():  # what the reference's runtest greps for
        assert got[k] == v


@pytest.mark.parametrize("algo", [0, 1, 2])
def test_example_trajectory(algo):
    """BASELINE.json configs[0]: examples/inputs 100 x 1000, seed 12345, -a 0/1/2"""
    want = SA_RUNS[f"a{algo}"]
    got = H.run_cli(H.example_input(), ["-s", "12345", "-a", str(algo)], mock=True)
    H.check_against(got, want, f"example a={algo}")
    assert got["score_launches"] < got["rearrangements"]  # proposals were batched


@pytest.mark.parametrize("k", [1, 3, 16])
def test_speculation_depth_does_not_change_the_result(k):
    """-K fixes the number of proposals scored per launch; the trajectory must not depend on it"""
    want = CLI_RUNS["test_treelength_7_thread_2@1"]
    got = H.run_cli(H.cli_input("test_treelength_7_thread_2"), want["flags"], mock=True, extra=["-K", str(k)])
    H.check_against(got, want, f"K={k}")
    if k == 1:
        assert got["trees_scored"] == want["rearrangements"] + 2 + 101 * round(float(want["start_temp"]) / 1e-5 + 1)


def test_replicates_are_independent_chains():
    """-R 3: chain r is the single-chain run with seed s + r (BASELINE.json configs[4] in small)"""
    p = subprocess.run([sys.executable, os.path.join(HERE, "run_host_search.py"), "--mock", "treelength.npz:tl5_thread_2_text",
                        "1", "1", "0", "3", "0", "3"], capture_output=True, text=True, timeout=900)  # 3 host threads
    assert p.returncode == 0, p.stderr[-2000:]
    multi = json.loads(p.stdout.strip().splitlines()[-1])
    for r in range(3):
        q = subprocess.run([sys.executable, os.path.join(HERE, "run_host_search.py"), "--mock",
                            "treelength.npz:tl5_thread_2_text", str(1 + r), "1"], capture_output=True, text=True, timeout=900)
        assert q.returncode == 0, q.stderr[-2000:]
        single = json.loads(q.stdout.strip().splitlines()[-1])
        a, b = multi["results"][r], single["results"][0]
        for k in ("final_length", "rearrangements", "topologies", "initial_length", "start_temperature"):
            assert a[k] == b[k], (r, k)
        assert multi["outtrees"][r] == single["outtrees"][0]
        if r == 0:
            assert multi["tsv_md5"] == single["tsv_md5"]


def test_chain_groups_do_not_change_the_result():
    """9 chains, 4 host threads, 2 groups, each with its own device context and driver thread and
    each worked as two alternating halves (GetSolnBatched); chain r is still the single-chain run
    with seed s + r"""
    p = subprocess.run([sys.executable, os.path.join(HERE, "run_host_search.py"), "--mock", "treelength.npz:tl5_thread_2_text",
                        "7", "1", "0", "9", "0", "4", "2"], capture_output=True, text=True, timeout=900)
    assert p.returncode == 0, p.stderr[-2000:]
    multi = json.loads(p.stdout.strip().splitlines()[-1])
    assert multi["results"][0]["groups"] == 2
    for r in (3, 4, 8):
        q = subprocess.run([sys.executable, os.path.join(HERE, "run_host_search.py"), "--mock",
                            "treelength.npz:tl5_thread_2_text", str(7 + r), "1"], capture_output=True, text=True, timeout=900)
        assert q.returncode == 0, q.stderr[-2000:]
        single = json.loads(q.stdout.strip().splitlines()[-1])
        a, b = multi["results"][r], single["results"][0]
        for k in ("final_length", "rearrangements", "topologies", "initial_length", "start_temperature"):
            assert a[k] == b[k], (r, k)
        assert multi["outtrees"][r] == single["outtrees"][0]


def test_matchange_on_the_device_route(monkeypatch):
    """LVB_GPU_MATCHANGE=1: constchar + MinimumTreeLength come from lvb_gpu_column_stats
    (matchange_device); same run as with the host scans"""
    monkeypatch.setenv("LVB_GPU_MATCHANGE", "1")
    for case in ("test_treelength_5_thread_2", "test_treelength_4@2"):
        want = CLI_RUNS[case]
        got = H.run_cli(H.cli_input(case.split("@")[0]), want["flags"], mock=True)
        H.check_against(got, want, case)
