"""Test infrastructure: host-layer operators (lvb_b200/host) against the reference itself
(oracle/_ref), same seeds, same call order.  Own process per check because the reference sizes
static scratch by the first alignment it sees.  usage: live_host_checks.py ops n m seed"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, HERE)
import oracles  # noqa: E402
from lvb_b200 import host  # noqa: E402


def same(a, b, what):
    pa, la, ra, _, da = a.get()
    pb, lb, rb, db = b.get()
    n = a.msa.n
    assert np.array_equal(pa, pb), (what, "parent", np.nonzero(pa != pb))
    assert np.array_equal(la, lb), (what, "left")
    assert np.array_equal(ra, rb), (what, "right")
    assert np.array_equal(da[n:], db[n:]), (what, "dirty", np.nonzero(da != db))
    assert a.root == b.root, (what, "root")


def ops(n, m, seed, steps=3000):
    rng = np.random.default_rng(seed)
    amb = np.frombuffer(b"ACGTACGTACGTUYRWSKMBDHVNX?-", dtype=np.uint8)
    mat = amb[rng.integers(0, len(amb), size=(n, m))]
    mat[:, rng.random(m) < 0.1] = ord("A")  # constant columns for matchange
    rows = [bytes(r) for r in mat]
    ref = oracles.RefMSA.from_rows(rows)
    h = host.HostMSA.from_rows(rows)
    assert (h.n, h.m, h.original_m, h.nwords, h.bytes, h.B, h.min_len) == (
        ref.n, ref.m, ref.original_m, ref.nwords, ref.bytes, ref.B, ref.min_len)
    assert h.rows_text() == ref.rows_text()
    assert np.array_equal(h.enc(), ref.enc())
    L = oracles.ref()
    L.ref_rinit(seed + 5)
    h.rinit(seed + 5)
    rc, hc = ref.tree().random(), h.tree().random()
    same(rc, hc, "PullRandomTree")
    rc.getplen()
    hc.mark_clean()
    rp, hp = ref.tree(), h.tree()
    for i in range(steps):
        kind = ("nni", "spr", "tbr")[i % 3]
        rp.mutate(kind, rc)
        hp.mutate(kind, hc)
        same(rp, hp, (i, kind))
        assert L.ref_uni() == h.uni(), (i, "stream position")
        if i % 4 == 0:  # accept: the proposal becomes the current tree
            rp.getplen()
            hp.mark_clean()
            rc, rp, hc, hp = rp, rc, hp, hc
        if i % 50 == 49:
            rc.arbreroot()
            newroot, path = hc.arbreroot()
            same(rc, hc, (i, "arbreroot"))
            # the rows that change under the new root are exactly the internal nodes on the path
            assert all(v >= n for v in path)
            rc.getplen()
            hc.mark_clean()
        if i % 97 == 0 and n > 5:
            p = int(rng.integers(n, 2 * n - 3))
            left = bool(rng.integers(0, 2))
            rp.mutate_deterministic(rc, p, left)
            hp.mutate_deterministic(hc, p, left)
            same(rp, hp, (i, "mutate_deterministic"))
    # a second random tree continues the same stream
    same(ref.tree().random(), h.tree().random(), "PullRandomTree again")


def stack(n, m, seed, count=400):
    """duplicate detection of the tree stack.  One sequence of trees — proposals, the same tree
    again, the same topology re-rooted, trees pushed long before — goes into a fresh stack on both sides:
    CompareTreeToTreestack (Treestack.c:231-306) and CompareHashTreeToHashstack (Hash.cpp:49-90) of the reference
    against lvb_b200/host/treestack.cpp; every push must give the same answer (1 = new, 0 = duplicate)."""
    import ctypes as C

    rng = np.random.default_rng(seed)
    mat = np.frombuffer(b"ACGT", dtype=np.uint8)[rng.integers(0, 4, size=(n, m))]
    rows = [bytes(r) for r in mat]
    ref = oracles.RefMSA.from_rows(rows)
    h = host.HostMSA.from_rows(rows)
    L = oracles.ref()
    L.ref_rinit(seed + 9)
    h.rinit(seed + 9)
    rc, hc = ref.tree().random(), h.tree().random()
    rc.getplen()
    hc.mark_clean()
    rts, hts, kinds = [], [], []

    def keep(what):
        rts.append(ref.tree().copy_from(rc))
        hts.append(h.tree().copy_from(hc))
        kinds.append(what)

    keep("start")
    for i in range(count):
        r = int(rng.integers(0, 10))
        if r < 5:  # a rearrangement becomes the current tree
            kind = ("nni", "spr", "tbr")[i % 3]
            rp, hp = ref.tree().mutate(kind, rc), h.tree().mutate(kind, hc)
            rp.getplen()
            hp.mark_clean()
            rc, hc = rp, hp
            keep(kind)
        elif r < 7:  # the same unrooted topology seen from another root
            rc.arbreroot()
            hc.arbreroot()
            assert rc.root == hc.root
            keep("reroot")
        elif r < 8:  # exactly the same tree again
            keep("again")
        else:  # a tree that went in earlier
            j = int(rng.integers(0, len(rts)))
            rts.append(rts[j]); hts.append(hts[j]); kinds.append(("old", j))
    k = len(rts)
    roots_r = (C.c_long * k)(*[t.root for t in rts])
    roots_h = (C.c_long * k)(*[t.root for t in hts])
    assert list(roots_r) == list(roots_h)
    ptr_r = (C.c_void_p * k)(*[t.t for t in rts])
    ptr_h = (C.c_void_p * k)(*[t.t for t in hts])
    L.ref_treestack_sequence.restype = C.c_long
    L.ref_treestack_sequence.argtypes = [C.c_void_p, C.c_long, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p]
    HL = host.load_host_library()
    HL.lvbh_treestack_sequence.restype = C.c_long
    HL.lvbh_treestack_sequence.argtypes = [C.c_void_p, C.c_long, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p]
    for use_hash in (0, 1):
        res_r, res_h = (C.c_long * k)(), (C.c_long * k)()
        size_r = L.ref_treestack_sequence(ref.h, k, ptr_r, roots_r, use_hash, res_r)
        size_h = HL.lvbh_treestack_sequence(h.h, k, ptr_h, roots_h, use_hash, res_h)
        bad = [(i, kinds[i], res_r[i], res_h[i]) for i in range(k) if res_r[i] != res_h[i]]
        assert not bad, (use_hash, bad[:5])
        assert size_r == size_h and 1 < size_r < k, (use_hash, size_r, size_h, k)
        assert any(res_r[i] == 0 for i in range(k) if kinds[i] == "reroot")  # a re-rooted tree IS a duplicate


if __name__ == "__main__":
    globals()[sys.argv[1]](*[int(a) for a in sys.argv[2:]])
    print("ok")
