"""Test infrastructure: checks that call the real reference (oracle/_ref).  Each runs in its
own process (see test_oracle.py) because the reference sizes static scratch by the first
alignment it sees.  Usage: python live_checks.py <check> [args...]"""
import ctypes as C
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import oracles  # noqa: E402


def text_rows(mat):
    return [bytes(r) for r in np.asarray(mat, dtype=np.uint8)]


def random_alignment(rng, n, m, ambiguity=0.05, constant=0.1):
    alphabet = np.frombuffer(b"ACGT", dtype=np.uint8)
    amb = np.frombuffer(b"UYRWSKMBDHVNX?-", dtype=np.uint8)
    mat = alphabet[rng.integers(0, 4, size=(n, m))]
    mask = rng.random((n, m)) < ambiguity
    mat[mask] = amb[rng.integers(0, len(amb), size=int(mask.sum()))]
    const_cols = rng.random(m) < constant
    mat[:, const_cols] = mat[0, const_cols]
    return mat


def oracle_vs_reference(n, m, seed):
    rng = np.random.default_rng(seed)
    mat = random_alignment(rng, n, m)
    rows = text_rows(mat)
    msa = oracles.RefMSA.from_rows(rows)
    # matchange: constant columns cut (DataOperations.c:272-405)
    L = oracles.oracle()
    keep = np.zeros(m, dtype=np.uint8)
    arr = (C.c_char_p * n)(*rows)
    kept = L.oracle_variable_columns(n, m, arr, keep.ctypes.data)
    assert kept == msa.m
    filt = text_rows(mat[:, keep.astype(bool)])
    assert filt == msa.rows_text()
    arrf = (C.c_char_p * n)(*filt)
    assert L.oracle_min_tree_length(n, kept, arrf) == msa.min_len
    enc = oracles.oracle_pack(filt)
    assert np.array_equal(enc, msa.enc())
    # trees: full pass then a run of proposals, incremental on both sides
    oracles.ref().ref_rinit(100 + seed)
    cur = msa.tree().random()
    want = cur.getplen()
    _, left, right, _, _ = cur.get()
    length, orows, ochg = oracles.oracle_getplen(enc, left, right, cur.root)
    assert length == want
    assert np.array_equal(orows, cur.rows())
    prop = msa.tree()
    for i in range(30):
        prop.mutate(("nni", "spr", "tbr")[i % 3], cur)
        _, left, right, _, dirty = prop.get()
        want = prop.getplen()
        rows2 = orows.copy()
        rows2[dirty.astype(bool), 0] = 0
        length, rows2, chg2 = oracles.oracle_getplen(enc, left, right, prop.root, rows2, ochg.copy())
        assert length == want, i
        assert np.array_equal(rows2, prop.rows())
        if i % 2 == 0:  # accept
            cur, prop = prop, cur
            orows, ochg = rows2, chg2


def threads_vs_sliced():
    golden_dir = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
    g = np.load(os.path.join(golden_dir, "treelength.npz"))
    rows = text_rows(g["tl6_thread_2_text"])
    msa = oracles.RefMSA.from_rows(rows, threads=2)
    assert (msa.n_threads, msa.slice) == (2, 35)  # the one reference test that takes the OpenMP path
    oracles.ref().ref_rinit(7)
    t = msa.tree().random()
    want = t.getplen()
    _, left, right, _, _ = t.get()
    for s in (1, 2):
        length, orows, _ = oracles.oracle_getplen(msa.enc(), left, right, t.root, slices=s)
        assert length == want
        assert np.array_equal(orows, t.rows())


if __name__ == "__main__":
    fn = globals()[sys.argv[1]]
    fn(*[int(a) for a in sys.argv[2:]])
    print("ok")
