"""Test infrastructure shared by test_host_search.py (CPU, mock device) and test_gpu_search.py
(real device): run the `lvb_gpu` binary the way the reference's own executable tests run `lvb`
(test/src/COMMON/*/runtest: a directory holding `infile`, regexes on stdout) and collect what
tests/golden/cli_runs.json / sa_runs.json record for the unmodified reference."""
import hashlib
import os
import re
import subprocess
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
CLI = os.path.join(ROOT, "lvb_b200", "bin", "lvb_gpu")


def mock_library() -> str:
    import run_host_search

    return run_host_search.build_mock()


def md5_file(path: str) -> str:
    return hashlib.md5(open(path, "rb").read()).hexdigest()


def run_cli(infile_bytes: bytes, flags, mock: bool, timeout: int = 900, extra=()):
    """returns dict(score, topologies, rearrangements, start_temp, tsv_md5, outtree_md5, stdout, rc, ...)"""
    env = dict(os.environ)
    if mock:
        env["LD_PRELOAD"] = mock_library()
    with tempfile.TemporaryDirectory() as td:
        open(os.path.join(td, "infile"), "wb").write(infile_bytes)
        p = subprocess.run([CLI, *flags, "-v", *extra], cwd=td, capture_output=True, text=True, env=env, timeout=timeout)
        o = p.stdout
        out = dict(rc=p.returncode, stdout=o, stderr=p.stderr)
        if p.returncode != 0:
            return out

        def num(pat, cast=int):
            m = re.search(pat, o)
            return cast(m.group(1)) if m else None

        out.update(
            score=num(r"Tree score: +(\d+)\n"),
            topologies=num(r"Topologies recovered: +(\d+)\n"),
            rearrangements=num(r"Rearrangements evaluated: +(\d+)\n"),
            start_temp=num(r"Starting Temperature: +([0-9.]+)", str),
            trees_scored=num(r"trees scored: (\d+)"),
            score_launches=num(r"trees scored: \d+ in (\d+) launches"),
            kernel_launches=num(r"kernel launches: (\d+)"),
            search_seconds=num(r"Wall time of the search: +([0-9.]+) s", float),
            tsv_md5=md5_file(os.path.join(td, "changeAccepted.tsv")),
            outtree_md5=md5_file(os.path.join(td, "outtree")),
            init_len=int(open(os.path.join(td, "sum")).read().splitlines()[1].split("\t")[2]),
        )
        return out


def check_against(got: dict, want: dict, what: str = "") -> None:
    assert got["rc"] == 0, (what, got.get("stdout", "")[-800:], got.get("stderr", "")[-800:])
    for k in ("score", "topologies", "rearrangements", "start_temp", "init_len", "tsv_md5", "outtree_md5"):
        assert got[k] == want[k], (what, k, got[k], want[k])


def cli_input(case_dir: str) -> bytes:
    return np.load(os.path.join(HERE, "golden", "cli_inputs.npz"))[case_dir].tobytes()


def example_input() -> bytes:
    return np.load(os.path.join(HERE, "golden", "example_phy.npz"))["raw"].tobytes()


def config2_input() -> bytes:
    import sys

    sys.path.insert(0, ROOT)
    from lvb_b200 import synthetic

    with tempfile.TemporaryDirectory() as td:
        path = os.path.join(td, "c2.phy")
        synthetic.write_phylip(path, synthetic.survey_alignment(100, 10000, 1, 0.05))
        raw = open(path, "rb").read()
    assert hashlib.md5(raw).hexdigest() == "a11fa9393aaf6c6b8150b4491e8a4d90"
    return raw
