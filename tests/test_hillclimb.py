"""deterministic_hillclimb (Solve.c:64-142): the host loop over
lvb_gpu_nni_neighbourhood against the UNMODIFIED reference's own function from the same random
start tree — final length, iteration count, trees on the stack and the Newick output (md5).
The reference values were recorded from oracle/_ref (tests/run_hillclimb.py --ref) and are
re-derived live whenever oracle/_ref is present."""
import json
import os
import subprocess
import sys

import pytest

import oracles

HERE = os.path.dirname(os.path.abspath(__file__))

RECORDED = {  # run_hillclimb.py --ref <key> <seed>
    ("treelength.npz:tl5_thread_2_text", 7): dict(length=1846, iterations=42, trees=1, outtree_md5="ed25febd0c3c0f14c5be241080e197de"),
    ("treelength.npz:tl6_thread_3_text", 7): dict(length=297, iterations=432, trees=9, outtree_md5="33606225eb95171a4acdaf40c1b5b78e"),
    ("treelength.npz:tl4_text", 7): dict(length=9, iterations=102, trees=74, outtree_md5="420cefec5df81fc5ca0154ce67454a40"),
    ("example.npz:text", 12345): dict(length=4232, iterations=4074, trees=6, outtree_md5="71b4c877a25a72757028ba130d7e7fa8"),
}


def run(mode, key, seed):
    cmd = [sys.executable, os.path.join(HERE, "run_hillclimb.py")] + ([mode] if mode else []) + [key, str(seed)]
    p = subprocess.run(cmd, capture_output=True, text=True, timeout=900)
    assert p.returncode == 0, p.stdout[-1000:] + p.stderr[-2000:]
    return json.loads(p.stdout.strip().splitlines()[-1])


def check(got, key, seed):
    want = RECORDED[(key, seed)]
    if oracles.have_ref():
        live = run("--ref", key, seed)
        assert {k: live[k] for k in want} == want, "recorded reference values are stale"
    for k, v in want.items():
        assert got[k] == v, (k, got, want)
    # one neighbourhood call per accepted move + one per start, never one per candidate
    assert got["neighbourhood_calls"] < got["iterations"]


@pytest.mark.parametrize("key,seed", [k for k in RECORDED if not k[0].startswith("example")])
def test_hillclimb_host_logic_on_the_mock_device(key, seed):
    check(run("--mock", key, seed), key, seed)


@pytest.mark.gpu
@pytest.mark.parametrize("key,seed", list(RECORDED))
def test_hillclimb_on_the_gpu(key, seed):
    check(run(None, key, seed), key, seed)
