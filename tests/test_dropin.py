"""The drop-in boundary demonstrated: the UNMODIFIED reference — its own Main,
Solve, StartingTemperature, TreeOperations, Treestack ... objects — linked with
integration/TreeEvaluation_gpu.c in place of TreeEvaluation.o (oracle/Makefile target
_ref/lvb_ref_gpu).  Its getplen() has the reference's signature (src/LVB.h:187), is called by the
reference's own callers (Solve.c:200,300; StartingTemperature.c:98,128) and gets every length from
the lvb_gpu_* C-ABI.  The whole search must come out byte-identical to the reference binary's.

CPU: the oracle-backed mock device (tests/mock_gpu.c) is preloaded, which checks the protocol
(tags, treecopy provenance, dirty lists).  GPU: the real liblvbgpu.so.
"""
import os
import subprocess
import tempfile

import pytest

import cli_harness as H

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF_GPU = os.path.join(ROOT, "oracle", "_ref", "lvb_ref_gpu")

# SURVEY.md 8(c), `lvb -i <example> -s 12345 -a 1`: outtree / changeAccepted.tsv of the unmodified reference
WANT_OUTTREE = "841b958a23fb63d88c79feb95e3fcfd6"
WANT_TSV = "859cf86578b87df2b8ab74861b17fbf4"

needs_binary = pytest.mark.skipif(not os.path.exists(REF_GPU), reason="oracle/_ref/lvb_ref_gpu not built on this box")


def run_dropin(flags, mock: bool, raw: bytes = None):
    env = dict(os.environ, LVB_GPU_REPORT="1")
    if mock:
        env["LD_PRELOAD"] = H.mock_library()
    with tempfile.TemporaryDirectory() as td:
        open(os.path.join(td, "infile"), "wb").write(raw if raw is not None else H.example_input())
        p = subprocess.run([REF_GPU, *flags, "-v"], cwd=td, capture_output=True, text=True, env=env, timeout=1500)
        assert p.returncode == 0, (p.stdout[-800:], p.stderr[-800:])
        launches = [int(l.split(":")[1]) for l in p.stderr.splitlines() if l.startswith("lvb_gpu kernel launches:")]
        return dict(stdout=p.stdout, outtree=H.md5_file(os.path.join(td, "outtree")),
                    tsv=H.md5_file(os.path.join(td, "changeAccepted.tsv")), launches=launches[0] if launches else 0)


@needs_binary
def test_reference_with_replaced_tree_evaluation_mock_device():
    got = run_dropin(["-s", "12345", "-a", "1", "-p", "1"], mock=True)
    assert (got["outtree"], got["tsv"]) == (WANT_OUTTREE, WANT_TSV)
    assert "Tree score:               4199" in got["stdout"]
    assert got["launches"] > 0


@needs_binary
def test_dropin_ignores_thread_plan_mock_device():
    """-p 2 makes the reference's callers allocate the OpenMP scratch (MemoryOperations.c:120-164);
    the replacement accepts and ignores it — same run"""
    got = run_dropin(["-s", "12345", "-a", "1", "-p", "2"], mock=True)
    assert (got["outtree"], got["tsv"]) == (WANT_OUTTREE, WANT_TSV)


@pytest.mark.gpu
@needs_binary
def test_reference_with_replaced_tree_evaluation_on_gpu():
    got = run_dropin(["-s", "12345", "-a", "1", "-p", "1"], mock=False)
    assert (got["outtree"], got["tsv"]) == (WANT_OUTTREE, WANT_TSV)
    assert got["launches"] > 0


@pytest.mark.gpu
@needs_binary
def test_dropin_reference_test_input_on_gpu():
    """one of the reference's own executable tests (ambiguity codes) through the replaced getplen"""
    import json

    want = json.load(open(os.path.join(ROOT, "tests", "golden", "cli_runs.json")))["test_treelength_4@2"]
    got = run_dropin(list(want["flags"]), mock=False, raw=H.cli_input("test_treelength_4"))
    assert got["outtree"] == want["outtree_md5"] and got["tsv"] == want["tsv_md5"]
