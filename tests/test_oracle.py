"""CPU: pin the oracle (oracle/fitch_oracle.c) against the reference's golden vectors
(tests/golden, made by tests/make_golden.py from the unmodified reference) and, where
oracle/_ref is built, against the reference itself on fresh random inputs."""
import ctypes as C
import os

import numpy as np
import pytest

import oracles


def load(golden_dir, name):
    return np.load(os.path.join(golden_dir, name))


def text_rows(mat):
    return [bytes(r) for r in np.asarray(mat, dtype=np.uint8)]


def test_words_per_row():
    L = oracles.oracle()
    for m, w in [(1, 1), (15, 1), (16, 1), (17, 2), (984, 62), (1000, 63), (10000, 625), (1000000, 62500)]:
        assert L.oracle_words_per_row(m) == w


def test_pack_example_known_words(golden_dir):
    g = load(golden_dir, "example.npz")
    enc = oracles.oracle_pack(text_rows(g["text"]))
    assert np.array_equal(enc, g["enc"])
    # check values measured on the unmodified reference
    assert int(enc[0, 0]) == 0x2428844484111288
    assert int(enc[0, -1]) == 0xFFFFFFFF28811818


def test_pack_and_minlen_treelength_inputs(golden_dir):
    g = load(golden_dir, "treelength.npz")
    L = oracles.oracle()
    for key in ["tl1", "tl2", "tl3", "tl4", "tl5_thread_2", "tl6_thread_2", "tl6_thread_3", "tl7_thread_2"]:
        rows = text_rows(g[f"{key}_text"])
        assert np.array_equal(oracles.oracle_pack(rows), g[f"{key}_enc"]), key
        arr = (C.c_char_p * len(rows))(*rows)
        assert L.oracle_min_tree_length(len(rows), len(rows[0]), arr) == int(g[f"{key}_min_len"]), key


def test_pack_rejects_bad_symbol():
    with pytest.raises(ValueError):
        oracles.oracle_pack([b"ACGTO", b"ACGTA", b"ACGTA", b"ACGTA", b"ACGTA"])
    with pytest.raises(ValueError):
        oracles.oracle_pack([b"acgt", b"ACGT", b"ACGT", b"ACGT", b"ACGT"])


def test_fitch_word_truth_table():
    """every pair of 4-bit state sets, in every nibble position"""
    L = oracles.oracle()
    for k in (0, 7, 15):
        for x in range(1, 16):
            for y in range(1, 16):
                ones = 0xFFFFFFFFFFFFFFFF & ~(0xF << (4 * k))
                out = C.c_uint64(0)
                ch = L.oracle_fitch_word(ones | (x << (4 * k)), ones | (y << (4 * k)), C.byref(out))
                want = (x & y) if (x & y) else (x | y)
                assert ch == (0 if (x & y) else 1)
                assert (out.value >> (4 * k)) & 0xF == want
                assert out.value | (0xF << (4 * k)) == 0xFFFFFFFFFFFFFFFF


def test_getplen_known_answers(golden_dir):
    """SURVEY.md §8(c) chain: 11808, nni 11834, spr 11835, tbr 11827, reroot 11827"""
    g = load(golden_dir, "example.npz")
    enc = g["enc"]
    for step, want in [("random", 11808), ("nni", 11834), ("spr", 11835), ("tbr", 11827), ("reroot", 11827)]:
        assert int(g[f"{step}_length"]) == want
        length, rows, changes = oracles.oracle_getplen(enc, g[f"{step}_left"], g[f"{step}_right"], int(g[f"{step}_root"]))
        assert length == want
        assert np.array_equal(rows, g[f"{step}_rows_after"])
        n = enc.shape[0]
        assert np.array_equal(changes[n:], g[f"{step}_changes_after"][n:])


def test_getplen_incremental_matches_reference_dirty_protocol(golden_dir):
    """start from the evaluated random tree, apply the recorded NNI topology with the recorded
    dirty flags: only dirty rows are recomputed and the length matches"""
    g = load(golden_dir, "example.npz")
    enc = g["enc"]
    n = enc.shape[0]
    rows = g["random_rows_after"].copy()
    changes = g["random_changes_after"].copy()
    dirty = g["nni_dirty"].astype(bool)
    assert 0 < dirty[n:].sum() < n - 3
    rows[dirty, 0] = 0
    length, rows, changes = oracles.oracle_getplen(enc, g["nni_left"], g["nni_right"], int(g["nni_root"]), rows, changes)
    assert length == 11834
    assert np.array_equal(rows, g["nni_rows_after"])
    assert np.array_equal(changes[n:], g["nni_changes_after"][n:])


def test_getplen_sliced_equals_serial(golden_dir):
    g = load(golden_dir, "example.npz")
    enc = g["enc"]
    for s in (1, 2, 3, 8, 31, 62):
        length, rows, changes = oracles.oracle_getplen(enc, g["tbr_left"], g["tbr_right"], int(g["tbr_root"]), slices=s)
        assert length == 11827
        assert np.array_equal(rows, g["tbr_rows_after"])


def test_getplen_proposal_series(golden_dir):
    g = load(golden_dir, "proposals.npz")
    enc = load(golden_dir, "example.npz")["enc"]
    for i in range(0, len(g["length"]), 7):
        length, _, _ = oracles.oracle_getplen(enc, g["left"][i], g["right"][i], int(g["root"][i]))
        assert length == int(g["length"][i])


def test_getplen_treelength_inputs(golden_dir):
    g = load(golden_dir, "treelength.npz")
    for key in ["tl1", "tl2", "tl3", "tl4", "tl5_thread_2", "tl6_thread_2", "tl6_thread_3", "tl7_thread_2"]:
        enc = g[f"{key}_enc"]
        lens = g[f"{key}_length"]
        for i in range(len(lens)):
            length, _, _ = oracles.oracle_getplen(enc, g[f"{key}_left"][i], g[f"{key}_right"][i], int(g[f"{key}_root"][i]))
            assert length == int(lens[i]), (key, i)
        # no tree can be shorter than the score the reference's own test pins for this input
        assert lens.min() >= int(g[f"{key}_score"]), key


def test_sa_goldens_are_the_survey_runs(golden_dir):
    import json

    r = json.load(open(os.path.join(golden_dir, "sa_runs.json")))
    assert (r["a0"]["rearrangements"], r["a0"]["topologies"], r["a0"]["score"]) == (337893, 54, 4199)
    assert (r["a1"]["rearrangements"], r["a1"]["topologies"], r["a1"]["score"]) == (249794, 14, 4199)
    assert (r["a2"]["rearrangements"], r["a2"]["topologies"], r["a2"]["score"]) == (315290, 49, 4199)


# ---- live comparison with the reference build (skipped where oracle/_ref is absent) ----
# The reference keeps function-static scratch sized by the FIRST alignment it sees
# (TreeOperations.c:360, 585), so every live case runs in its own process.
needs_ref = pytest.mark.skipif(not oracles.have_ref(), reason="oracle/_ref not built")


def run_isolated(*args):
    import subprocess
    import sys

    p = subprocess.run([sys.executable, os.path.join(os.path.dirname(__file__), "live_checks.py"), *map(str, args)],
                       capture_output=True, text=True)
    assert p.returncode == 0, p.stdout[-2000:] + p.stderr[-2000:]


@needs_ref
@pytest.mark.parametrize("n,m,seed", [(5, 3, 1), (7, 16, 2), (12, 33, 3), (40, 1000, 4), (9, 2049, 5)])
def test_oracle_matches_reference_live(n, m, seed):
    run_isolated("oracle_vs_reference", n, m, seed)


@needs_ref
def test_reference_threads_agree_with_sliced_oracle():
    run_isolated("threads_vs_sliced")
