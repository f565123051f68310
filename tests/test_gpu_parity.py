"""GPU (-m gpu): the CUDA path, called through the C-ABI (include/lvb_gpu.h), against the
oracle and the reference's golden vectors.  Integer work: every comparison is bit-exact."""
import os

import numpy as np
import pytest

import oracles
from lvb_b200 import Engine, GpuError, Topology, synthetic

pytestmark = pytest.mark.gpu


def load(golden_dir, name):
    return np.load(os.path.join(golden_dir, name))


def all_rows(eng, tree, n):
    return np.stack([eng.download_row(tree, v) for v in range(2 * n - 3)])


def closure_dirty(left, right, root, cur_l, cur_r, n):
    """internal nodes whose children differ from the device's current tree, closed towards the root"""
    B = 2 * n - 3
    changed = np.nonzero((left != cur_l) | (right != cur_r))[0]
    parent = np.full(B, -1, dtype=np.int64)
    for v in range(n, B):
        parent[left[v]] = v
        parent[right[v]] = v
    dirty = set()
    for v in changed[changed >= n].tolist():
        while v >= n and v not in dirty:
            dirty.add(v)
            v = int(parent[v])
    return np.array(sorted(dirty), dtype=np.int64)


def test_known_answer_chain(engine, golden_dir):
    """11808 -> nni 11834 -> spr 11835 -> tbr 11827 -> reroot 11827, with
    every state row and every per-node change count equal to the reference's"""
    g = load(golden_dir, "example.npz")
    enc = g["enc"]
    n = enc.shape[0]
    engine.upload(enc, max_trees=2)
    t = engine.tree_create()
    assert engine.commit(t, Topology(g["random_left"], g["random_right"], 0)) == 11808
    assert np.array_equal(all_rows(engine, t, n), g["random_rows_after"])
    assert np.array_equal(engine.changes(t)[n:], g["random_changes_after"][n:])
    cur_l, cur_r = g["random_left"], g["random_right"]
    for step, want in [("nni", 11834), ("spr", 11835), ("tbr", 11827), ("reroot", 11827)]:
        left, right, root = g[f"{step}_left"], g[f"{step}_right"], int(g[f"{step}_root"])
        dirty = np.nonzero(g[f"{step}_dirty"])[0] if step != "reroot" else closure_dirty(left, right, root, cur_l, cur_r, n)
        topo = Topology(left, right, root, dirty)
        assert engine.score_batch([t], [topo]).tolist() == [want]
        assert np.array_equal(engine.changes(t)[n:], g[f"{'random' if step == 'nni' else prev}_changes_after"][n:])
        assert engine.commit(t, topo) == want
        assert np.array_equal(all_rows(engine, t, n), g[f"{step}_rows_after"]), step
        assert np.array_equal(engine.changes(t)[n:], g[f"{step}_changes_after"][n:]), step
        cur_l, cur_r, prev = left, right, step
    # full re-evaluation of the final tree gives the same state
    assert engine.commit(t, Topology(left, right, root)) == 11827
    assert np.array_equal(all_rows(engine, t, n), g["reroot_rows_after"])
    engine.tree_destroy(t)


def test_proposal_series(engine, golden_dir):
    """300 reference proposals: score against the device tree, commit the accepted ones"""
    g = load(golden_dir, "proposals.npz")
    enc = load(golden_dir, "example.npz")["enc"]
    n = enc.shape[0]
    engine.upload(enc, max_trees=1)
    t = engine.tree_create()
    cur_l, cur_r = g["start_left"].astype(np.int64), g["start_right"].astype(np.int64)
    engine.commit(t, Topology(cur_l, cur_r, int(g["start_root"])))
    pending, want = [], []
    for i in range(len(g["length"])):
        left, right, root = g["left"][i].astype(np.int64), g["right"][i].astype(np.int64), int(g["root"][i])
        ref_dirty = np.nonzero(g["dirty"][i])[0]
        dirty = ref_dirty if len(ref_dirty) < n - 3 else closure_dirty(left, right, root, cur_l, cur_r, n)
        topo = Topology(left, right, root, dirty)
        pending.append(topo)
        want.append(int(g["length"][i]))
        if g["accepted"][i] or i == len(g["length"]) - 1:
            # everything proposed since the last accept was derived from the same current tree:
            # one launch scores the whole run
            got = engine.score_batch([t] * len(pending), pending)
            assert got.tolist() == want, i
            pending, want = [], []
            if g["accepted"][i]:
                assert engine.commit(t, topo) == int(g["length"][i])
                cur_l, cur_r = left, right
    engine.tree_destroy(t)


@pytest.mark.parametrize("key", ["tl1", "tl2", "tl3", "tl4", "tl5_thread_2", "tl6_thread_2", "tl6_thread_3", "tl7_thread_2"])
def test_reference_test_inputs(engine, golden_dir, key):
    """alignments of the reference's test_treelength_* (ambiguity codes, gaps, 1-word rows)"""
    g = load(golden_dir, "treelength.npz")
    enc = g[f"{key}_enc"]
    n = enc.shape[0]
    engine.upload(enc, max_trees=1)
    t = engine.tree_create()
    lens = g[f"{key}_length"]
    for i in range(len(lens)):
        left, right = g[f"{key}_left"][i].astype(np.int64), g[f"{key}_right"][i].astype(np.int64)
        assert engine.commit(t, Topology(left, right, int(g[f"{key}_root"][i]))) == int(lens[i]), (key, i)
    engine.tree_destroy(t)


def test_upload_paths_and_device_pack(engine, golden_dir):
    g = load(golden_dir, "example.npz")
    enc = g["enc"]
    n = enc.shape[0]
    rows = [bytes(r) for r in g["text"]]
    topo = Topology(g["random_left"], g["random_right"], 0)
    engine.upload_rows([enc[i].copy() for i in range(n)], max_trees=1)  # Solve.c:516-519 layout
    t = engine.tree_create()
    assert engine.commit(t, topo) == 11808
    engine.pack(rows, max_trees=1)  # DNAToBinary on the device
    t = engine.tree_create()
    for v in range(n):
        assert np.array_equal(engine.download_row(t, v), enc[v])
    assert engine.commit(t, topo) == 11808
    with pytest.raises(GpuError, match="bad base symbol"):
        engine.pack([b"ACGTO"] + [b"ACGTA"] * 4)


def test_pack_all_symbols(engine):
    rng = np.random.default_rng(11)
    sym = np.frombuffer(b"ACGTUYRWSKMBDHVNX?-", dtype=np.uint8)
    for n, m in [(5, 1), (6, 15), (7, 16), (8, 17), (9, 1000)]:
        mat = sym[rng.integers(0, len(sym), size=(n, m))]
        rows = [bytes(r) for r in mat]
        engine.pack(rows, max_trees=1)
        t = engine.tree_create()
        want = oracles.oracle_pack(rows)
        for v in range(n):
            assert np.array_equal(engine.download_row(t, v), want[v])


def test_site_blocks_sum_to_whole(golden_dir):
    """site-block sharding (TreeEvaluation.c:95-97 decomposition): partial lengths of disjoint
    word blocks add up to the whole, for full passes and for scored candidates"""
    g = load(golden_dir, "example.npz")
    enc = g["enc"]
    n, nwords = enc.shape
    topo = Topology(g["random_left"], g["random_right"], 0)
    cand = Topology(g["nni_left"], g["nni_right"], 0, np.nonzero(g["nni_dirty"])[0])
    for blocks in (2, 3, 8):
        bounds = [nwords * b // blocks for b in range(blocks + 1)]
        full, scored = 0, 0
        for b in range(blocks):
            e = Engine(0)
            e.upload(enc, word_begin=bounds[b], word_count=bounds[b + 1] - bounds[b], max_trees=1)
            t = e.tree_create()
            full += e.commit(t, topo)
            scored += int(e.score_batch([t], [cand])[0])
            assert np.array_equal(e.download_row(t, n + 3), g["random_rows_after"][n + 3, bounds[b]:bounds[b + 1]])
            e.close()
        assert (full, scored) == (11808, 11834)


def test_replicate_batches(engine):
    """several independent trees in one launch (SA replicates): commit_batch + score_batch"""
    n, m, R = 40, 3000, 6
    st = synthetic.states(n, m, seed=5, p=0.08)
    enc = synthetic.pack_states(st)
    engine.upload(enc, max_trees=R)
    trees = [engine.tree_create() for _ in range(R)]
    topos = [Topology(*synthetic.random_topology(n, seed=100 + r)) for r in range(R)]
    got = engine.commit_batch(trees, topos)
    want = [oracles.oracle_getplen(enc, tp.left, tp.right, tp.root) for tp in topos]
    assert got.tolist() == [w[0] for w in want]
    for r in range(R):
        assert np.array_equal(engine.changes(trees[r])[n:], want[r][2][n:])
    # device-to-device copy keeps rows and counts
    engine.tree_copy(trees[0], trees[3])
    assert np.array_equal(engine.changes(trees[0]), engine.changes(trees[3]))
    assert np.array_equal(engine.download_row(trees[0], n + 5), want[3][1][n + 5])
    with pytest.raises(GpuError, match="twice"):
        engine.commit_batch([trees[1], trees[1]], topos[:2])
    for t in trees:
        engine.tree_destroy(t)


def nni_neighbourhood(left, right, root, n):
    """every NNI candidate of a tree (mutate_deterministic, TreeOperations.c:105-158) with its
    dirty path; a plain-python restatement used only to make test inputs"""
    B = 2 * n - 3
    parent = np.full(B, -1, dtype=np.int64)
    for v in list(range(n, B)) + [root]:
        parent[left[v]] = v
        parent[right[v]] = v
    out = []
    for u in range(n, B):
        v = int(parent[u])
        a, b = int(left[u]), int(right[u])
        c = int(right[v]) if left[v] == u else int(left[v])
        for swap_a in (False, True):
            l2, r2 = left.copy(), right.copy()
            moved = a if swap_a else b
            if l2[v] == u:
                r2[v] = moved
            else:
                l2[v] = moved
            l2[u], r2[u] = (b, c) if swap_a else (a, c)
            dirty, w = [], u
            while w != root:
                dirty.append(w)
                w = int(parent[w])
            out.append(Topology(l2, r2, root, dirty))
    return out


@pytest.mark.parametrize("n,m", [(30, 500), (150, 20000)])
def test_nni_neighbourhood_batch(engine, n, m):
    """2(n-3) candidates of one tree in a single launch (deterministic_hillclimb's loop,
    Solve.c:64-142) against full oracle passes"""
    enc = synthetic.pack_states(synthetic.states(n, m, seed=9, p=0.05))
    left, right, root = synthetic.random_topology(n, seed=4)
    engine.upload(enc, max_trees=1)
    t = engine.tree_create()
    assert engine.commit(t, Topology(left, right, root)) == oracles.oracle_getplen(enc, left, right, root)[0]
    cands = nni_neighbourhood(left, right, root, n)
    assert len(cands) == 2 * (n - 3)
    got = engine.score_batch([t] * len(cands), cands)
    step = max(1, len(cands) // 40)
    for k in range(0, len(cands), step):
        assert got[k] == oracles.oracle_getplen(enc, cands[k].left, cands[k].right, root)[0], k
    engine.tree_destroy(t)


@pytest.mark.parametrize("n,m,seed", [(5, 40, 1), (6, 33, 2), (30, 500, 3), (150, 20000, 4), (400, 3000, 5)])
def test_nni_neighbourhood_call(engine, n, m, seed):
    """lvb_gpu_nni_neighbourhood (up-pass + four-row kernel) == one incremental getplen per
    mutate_deterministic candidate (Solve.c:100-111), and == full oracle passes on a sample"""
    enc = synthetic.pack_states(synthetic.states(n, m, seed=seed, p=0.08))
    left, right, root = synthetic.random_topology(n, seed=seed + 10)
    engine.upload(enc, max_trees=2)
    t, s = engine.tree_create(), engine.tree_create()
    topo = Topology(left, right, root)
    base = engine.commit(t, topo)
    cands = nni_neighbourhood(left, right, root, n)
    want = engine.score_batch([t] * len(cands), cands)
    got = engine.nni_neighbourhood(t, s, topo)
    assert np.array_equal(got, want)
    for k in range(0, len(cands), max(1, len(cands) // 12)):
        assert got[k] == oracles.oracle_getplen(enc, cands[k].left, cands[k].right, root)[0], k
    # the evaluated tree is untouched, the scratch tree is no longer an evaluated tree
    assert engine.score_batch([t], [Topology(left, right, root, [])])[0] == base
    with pytest.raises(GpuError, match="never been evaluated"):
        engine.score_batch([s], [Topology(left, right, root, [])])
    # after a rearrangement is committed the neighbourhood follows
    pick = cands[len(cands) // 2]
    engine.commit(t, pick)
    cands2 = nni_neighbourhood(pick.left, pick.right, root, n)
    got2 = engine.nni_neighbourhood(t, s, Topology(pick.left, pick.right, root))
    assert np.array_equal(got2, engine.score_batch([t] * len(cands2), cands2))
    engine.tree_destroy(t)
    engine.tree_destroy(s)


def _oracle_column_stats(rows):
    import ctypes as C

    L = oracles.oracle()
    n, m = len(rows), len(rows[0])
    arr = (C.c_char_p * n)(*rows)
    keep = np.zeros(m, dtype=np.uint8)
    kept = L.oracle_variable_columns(n, m, arr, keep.ctypes.data)
    cut = [bytes(np.frombuffer(r, dtype=np.uint8)[keep.astype(bool)]) for r in rows]
    arr2 = (C.c_char_p * n)(*cut)
    return keep, int(kept), int(L.oracle_min_tree_length(n, m, arr)), (int(L.oracle_min_tree_length(n, kept, arr2)) if kept else 0)


def test_column_stats_constchar_and_min_tree_length(engine, golden_dir):
    """lvb_gpu_column_stats == constchar + MinimumTreeLength (DataOperations.c:272-307, 53-105):
    the reference's test inputs (golden min_len values from the reference itself), every accepted
    symbol incl. ambiguity letters, > 5 states in a column, columns of only '-' (the -1 quirk)"""
    g = load(golden_dir, "treelength.npz")
    for key in [k[:-5] for k in g.files if k.endswith("_text")]:
        rows = [bytes(r) for r in g[key + "_text"]]
        keep, kept, all_len, kept_len = engine.column_stats(rows)
        okeep, okept, oall, okept_len = _oracle_column_stats(rows)
        assert np.array_equal(keep, okeep) and (kept, all_len, kept_len) == (okept, oall, okept_len), key
        if f"{key}_min_len" in g.files:
            # the reference's own MinimumTreeLength, taken after or before matchange depending on the fixture
            assert int(g[f"{key}_min_len"]) in (all_len, kept_len), key
    ex = load(golden_dir, "example.npz")
    rows = [bytes(r) for r in ex["text"]]  # already cut to 984 columns by the reference's matchange
    keep, kept, all_len, kept_len = engine.column_stats(rows)
    assert kept == 984 and keep.all() and all_len == kept_len == 1816  # min_len=1816
    rng = np.random.default_rng(3)
    sym = np.frombuffer(b"ACGTUYRWSKMBDHVNX?O-", dtype=np.uint8)
    n, m = 37, 1003
    t = sym[rng.integers(0, len(sym), size=(n, m))].copy()
    t[:, 5] = ord("-")           # no countable character: adds -1
    t[:, 6] = ord("A")           # constant
    t[:, 7] = ord("N"); t[3, 7] = ord("X")   # variable, but nothing countable
    t[:, 8] = sym[:1]; t[:6, 8] = sym[:6]    # exactly 6 distinct countable -> 5
    rows = [bytes(r) for r in t]
    got = engine.column_stats(rows)
    want = _oracle_column_stats(rows)
    assert np.array_equal(got[0], want[0]) and got[1:] == want[1:]
    assert got[0][5] == 0 and got[0][6] == 0 and got[0][7] == 1


def test_submit_and_wait_halves(engine):
    """lvb_gpu_score_batch_submit / lvb_gpu_batch_wait: several batches in flight, a commit of
    another tree in between, lengths identical to the synchronous call; misuse is loud"""
    n, m = 60, 4000
    enc = synthetic.pack_states(synthetic.states(n, m, seed=21, p=0.06))
    engine.upload(enc, max_trees=2)
    t0, t1 = engine.tree_create(), engine.tree_create()
    la, ra, roota = synthetic.random_topology(n, seed=1)
    lb, rb, rootb = synthetic.random_topology(n, seed=2)
    engine.commit(t0, Topology(la, ra, roota))
    engine.commit(t1, Topology(lb, rb, rootb))
    ca, cb = nni_neighbourhood(la, ra, roota, n), nni_neighbourhood(lb, rb, rootb, n)
    want_a = engine.score_batch([t0] * len(ca), ca)
    want_b = engine.score_batch([t1] * 7, cb[:7])
    ta = engine.score_batch_submit([t0] * len(ca), ca)
    tb = engine.score_batch_submit([t1] * 7, cb[:7])
    engine.commit(t1, cb[3], wait=False)  # enqueued behind both batches
    tc = engine.score_batch_submit([t0] * 5, ca[:5])
    assert np.array_equal(engine.batch_wait(tb), want_b)
    assert np.array_equal(engine.batch_wait(ta), want_a)
    assert np.array_equal(engine.batch_wait(tc), want_a[:5])
    with pytest.raises(GpuError, match="names no submitted batch"):
        engine.batch_wait(ta)
    tickets = [engine.score_batch_submit([t0], ca[:1]) for _ in range(8)]
    with pytest.raises(GpuError, match="waiting for lvb_gpu_batch_wait"):
        engine.score_batch_submit([t0], ca[:1])
    for tk in tickets:
        assert engine.batch_wait(tk)[0] == want_a[0]
    engine.tree_destroy(t0)
    engine.tree_destroy(t1)


def test_nni_neighbourhood_ambiguity_and_shapes(engine, golden_dir):
    """example alignment (padding: 984 sites) and caterpillar / balanced shapes with ambiguity codes"""
    g = np.load(os.path.join(golden_dir, "example.npz"))
    enc = g["enc"]
    n = enc.shape[0]
    engine.upload(enc, max_trees=2)
    t, s = engine.tree_create(), engine.tree_create()
    left, right, root = g["random_left"], g["random_right"], int(g["random_root"])
    assert engine.commit(t, Topology(left, right, root)) == 11808
    got = engine.nni_neighbourhood(t, s, Topology(left, right, root))
    cands = nni_neighbourhood(left, right, root, n)
    assert np.array_equal(got, engine.score_batch([t] * len(cands), cands))
    engine.tree_destroy(t)
    engine.tree_destroy(s)
    rng = np.random.default_rng(5)
    n, m = 40, 700
    st = rng.integers(1, 16, size=(n, m), dtype=np.uint8)  # every non-empty state set, ambiguity included
    enc = synthetic.pack_states(st)
    for shape in ("caterpillar", "balanced"):
        left, right, root = synthetic.shaped_topology(n, shape)
        engine.upload(enc, max_trees=2)
        t, s = engine.tree_create(), engine.tree_create()
        engine.commit(t, Topology(left, right, root))
        cands = nni_neighbourhood(left, right, root, n)
        got = engine.nni_neighbourhood(t, s, Topology(left, right, root))
        for k in range(len(cands)):
            assert got[k] == oracles.oracle_getplen(enc, cands[k].left, cands[k].right, root)[0], (shape, k)
        engine.tree_destroy(t)
        engine.tree_destroy(s)


@pytest.mark.parametrize("n,m,seed", [(5, 1, 1), (5, 16, 2), (6, 17, 3), (17, 4097, 4), (64, 8 * 16 * 256 + 5, 5), (500, 100000, 6)])
def test_full_pass_shapes(engine, n, m, seed):
    """ragged row widths (1 word, not a multiple of the 16-byte column, of the 128-byte pad,
    of the CTA width) up to BASELINE config 3's 500 x 100k"""
    st = synthetic.states(n, m, seed=seed, p=0.1)
    enc = synthetic.pack_states(st)
    left, right, root = synthetic.random_topology(n, seed=seed)
    engine.upload(enc, max_trees=1)
    t = engine.tree_create()
    want, rows, chg = oracles.oracle_getplen(enc, left, right, root)
    assert engine.commit(t, Topology(left, right, root)) == want
    assert np.array_equal(engine.changes(t)[n:], chg[n:])
    for v in {n, min((3 * n) // 2, 2 * n - 4), 2 * n - 4}:
        assert np.array_equal(engine.download_row(t, v), rows[v])
    engine.tree_destroy(t)


def test_caterpillar_and_balanced_trees(engine):
    """extreme shapes: dependency chain of n-3 nodes with no stack, and the deepest stack"""
    n, m = 129, 5000
    enc = synthetic.pack_states(synthetic.states(n, m, seed=3, p=0.1))
    B = 2 * n - 3
    left, right = np.full(B, -1, dtype=np.int64), np.full(B, -1, dtype=np.int64)
    left[0], right[0] = n, 1
    leaf = 2
    for v in range(n, B):
        if v + 1 < B:
            left[v], right[v] = v + 1, leaf
            leaf += 1
        else:
            left[v], right[v] = leaf, leaf + 1
    engine.upload(enc, max_trees=1)
    t = engine.tree_create()
    assert engine.commit(t, Topology(left, right, 0)) == oracles.oracle_getplen(enc, left, right, 0)[0]
    left, right = np.full(B, -1, dtype=np.int64), np.full(B, -1, dtype=np.int64)
    level, nxt = list(range(1, n)), n
    while len(level) > 2:
        new = []
        for k in range(0, len(level), 2):
            left[nxt], right[nxt] = level[k], level[k + 1]
            new.append(nxt)
            nxt += 1
        level = new
    left[0], right[0] = level
    want, rows, _ = oracles.oracle_getplen(enc, left, right, 0)
    assert engine.commit(t, Topology(left, right, 0)) == want
    assert np.array_equal(engine.download_row(t, B - 1), rows[B - 1])


def test_errors_are_loud(engine, golden_dir):
    g = load(golden_dir, "example.npz")
    engine.upload(g["enc"], max_trees=1)
    t = engine.tree_create()
    topo = Topology(g["random_left"], g["random_right"], 0)
    with pytest.raises(GpuError, match="never been evaluated"):
        engine.score_batch([t], [topo])
    with pytest.raises(GpuError, match="first commit"):
        engine.commit(t, Topology(g["random_left"], g["random_right"], 0, [150]))
    with pytest.raises(GpuError, match="in use"):
        engine.tree_create()
    with pytest.raises(GpuError, match="does not exist"):
        engine.score_batch([5], [topo])
    with pytest.raises(GpuError, match="not a leaf"):
        engine.commit(t, Topology(g["random_left"], g["random_right"], 120))


def test_full_size_2000x1M(engine):
    """BASELINE config 4 shape on one GPU (2.0 GB of state rows): full pass bit-exact against
    the oracle, root-invariance of the length, and incremental == full on the same tree"""
    n, nwords = 2000, 62500
    enc = synthetic.random_packed(n, nwords, seed=2)
    left, right, root = synthetic.random_topology(n, seed=8)
    engine.upload(enc, max_trees=1)
    t = engine.tree_create()
    want, rows, chg = oracles.oracle_getplen(enc, left, right, root)
    assert engine.commit(t, Topology(left, right, root)) == want
    assert np.array_equal(engine.changes(t)[n:], chg[n:])
    # EVERY state row (1 GB), not a sample: the rows at the boundaries of the dependent pieces of the one-launch
    # route are the ones a scheduling mistake would get wrong
    bad = [v for v in range(n, 2 * n - 3) if not np.array_equal(engine.download_row(t, v), rows[v])]
    assert not bad, bad[:10]
    # a candidate: NNI at some deep node, scored incrementally, then committed, then re-evaluated in full
    cands = nni_neighbourhood(left, right, root, n)
    pick = [cands[k] for k in (0, 1777, 3993)]
    got = engine.score_batch([t] * 3, pick)
    full = oracles.oracle_getplen(enc, pick[1].left, pick[1].right, root)[0]
    assert got[1] == full
    assert engine.commit(t, pick[1]) == full
    assert engine.commit(t, Topology(pick[1].left, pick[1].right, root)) == full
    engine.tree_destroy(t)


def test_nni_neighbourhood_2000x1M(engine):
    """all 3994 mutate_deterministic candidates of a 2000 x 1M tree in one call == the incremental
    path on a sample of them (which test_full_size_2000x1M ties to the oracle)"""
    n, nwords = 2000, 62500
    enc = synthetic.random_packed(n, nwords, seed=2)
    left, right, root = synthetic.random_topology(n, seed=8)
    engine.upload(enc, max_trees=2)
    t, s = engine.tree_create(), engine.tree_create()
    engine.commit(t, Topology(left, right, root))
    got = engine.nni_neighbourhood(t, s, Topology(left, right, root))
    cands = nni_neighbourhood(left, right, root, n)
    idx = list(range(0, len(cands), 97)) + [len(cands) - 1]
    want = engine.score_batch([t] * len(idx), [cands[k] for k in idx])
    assert np.array_equal(got[idx], want)
    engine.tree_destroy(t)
    engine.tree_destroy(s)


@pytest.mark.parametrize("env", [
    {},                                                        # default: streaming kernel, tapered bins
    {"LVB_GPU_PIECE_MIN": "4", "LVB_GPU_TARGET_WARPS": str(148 * 4096)},  # many small pieces: every level of the cut
    {"LVB_GPU_PIECE_MIN": "4", "LVB_GPU_STORE_STREAM": "64", "LVB_GPU_STREAM_LAG": "3"},  # lagged work order
    {"LVB_GPU_TAPER": "0", "LVB_GPU_STORE_STREAM": "128"},
    {"LVB_GPU_STORE_STREAM": "0", "LVB_GPU_STORE_PF": "7", "LVB_GPU_PIECE_MIN": "8"},  # cp.async ring kernel
    {"LVB_GPU_STORE_STREAM": "0", "LVB_GPU_STORE_PF": "0", "LVB_GPU_PIECE_MIN": "8"},  # register-pipelined kernel
    {"LVB_GPU_STORE_STREAM": "0", "LVB_GPU_LEVEL_LAUNCHES": "1", "LVB_GPU_PIECE_MIN": "8"},  # one launch per level
])
def test_dependent_piece_store_routes(monkeypatch, env):
    """a whole-tree evaluation is cut into dependent pieces (forests in level-0 bins, the crown above
    them) and run in ONE launch: every kernel route and every setting of the cut gives the oracle's
    length, per-node counts and rows — three passes in a row (the level counters run on from launch to
    launch), for two topologies on the same device tree"""
    for k, v in env.items():
        monkeypatch.setenv(k, v)
    n, m = 700, 40_000
    enc = synthetic.pack_states(synthetic.states(n, m, seed=11, p=0.1))
    eng = Engine(0)
    try:
        eng.upload(enc, max_trees=2)
        t = eng.tree_create()
        for seed in (21, 22):
            left, right, root = synthetic.random_topology(n, seed=seed)
            want, rows, chg = oracles.oracle_getplen(enc, left, right, root)
            assert eng.commit(t, Topology(left, right, root)) == want
            eng.rerun_last_batch()
            eng.rerun_last_batch()
            eng.synchronize()
            assert eng.commit(t, Topology(left, right, root)) == want, (env, seed)
            assert np.array_equal(eng.changes(t)[n:], chg[n:]), (env, seed)
            for v in list(range(n, 2 * n - 3, 37)) + [2 * n - 4]:
                assert np.array_equal(eng.download_row(t, v), rows[v]), (env, seed, v)
    finally:
        eng.close()


def _nni_of(left, right, root, n, u, swap_a):
    """one NNI rearrangement around internal node u with its dirty path (plain test-input code)"""
    B = 2 * n - 3
    parent = np.full(B, -1, dtype=np.int64)
    for v in list(range(n, B)) + [root]:
        parent[left[v]] = v
        parent[right[v]] = v
    v = int(parent[u])
    a, b = int(left[u]), int(right[u])
    c = int(right[v]) if left[v] == u else int(left[v])
    l2, r2 = left.copy(), right.copy()
    moved = a if swap_a else b
    if l2[v] == u:
        r2[v] = moved
    else:
        l2[v] = moved
    l2[u], r2[u] = (b, c) if swap_a else (a, c)
    dirty, w = [], u
    while w != root:
        dirty.append(w)
        w = int(parent[w])
    return l2, r2, dirty


@pytest.mark.parametrize("env", [{}, {"LVB_GPU_INLINE": "0"}, {"LVB_GPU_FUSE": "0", "LVB_GPU_DEEP": "0"}])
def test_score_keep_and_accept(monkeypatch, env):
    """lvb_gpu_score_keep / lvb_gpu_accept_kept (Solve.c:300-323 without the second evaluation): kept candidates
    score like the read-only ones; an accepted one IS the tree afterwards — lengths of further candidates, per-node
    counts and rows as the oracle has them after a commit; one that is not accepted leaves the tree alone.  Twelve
    rounds on two trees, so rows move between the banks in both directions; a classic commit, a tree copy and the
    NNI neighbourhood in between read the rows where they are."""
    for k, v in env.items():
        monkeypatch.setenv(k, v)
    n, m = 150, 20_000
    enc = synthetic.pack_states(synthetic.states(n, m, seed=5, p=0.1))
    rng = np.random.default_rng(17)
    eng = Engine(0)
    try:
        eng.upload(enc, max_trees=4)
        trees = [eng.tree_create(), eng.tree_create()]
        scratch = eng.tree_create()
        cur = []
        for i, t in enumerate(trees):
            left, right, root = synthetic.random_topology(n, seed=30 + i)
            want, rows, chg = oracles.oracle_getplen(enc, left, right, root)
            assert eng.commit(t, Topology(left, right, root)) == want
            cur.append([left, right, root, rows, chg, want])
        for rnd in range(12):
            topos, which, keep, expect = [], [], [], []
            for i, t in enumerate(trees):
                left, right, root, rows, chg, _ = cur[i]
                for j in range(3):
                    u = int(rng.integers(n, 2 * n - 3))
                    l2, r2, dirty = _nni_of(left, right, root, n, u, bool(rng.integers(0, 2)))
                    topos.append(Topology(l2, r2, root, dirty))
                    which.append(t)
                    keep.append(1 if j == 1 else 0)  # the middle candidate of each tree is kept
                    rr, cc = rows.copy(), chg.copy()
                    rr[dirty, 0] = 0
                    expect.append((oracles.oracle_getplen(enc, l2, r2, root, rr, cc), l2, r2, dirty))
            lens, ticket = eng.score_keep(which, topos, keep)
            assert lens.tolist() == [e[0][0] for e in expect], rnd
            assert lens.tolist() == eng.score_batch(which, topos).tolist()
            # tree 0 accepts its kept candidate in even rounds, tree 1 in rounds 1, 2 mod 3; the others stay
            acc = [k for k in (1, 4) if (rnd % 2 == 0 if k == 1 else rnd % 3 != 0)]
            if rnd == 5:
                acc = []
            eng.accept_kept(ticket, acc)
            for k in acc:
                i = k // 3
                (want, rows, chg), l2, r2, dirty = expect[k]
                cur[i] = [l2, r2, cur[i][2], rows, chg, want]
            for i, t in enumerate(trees):
                left, right, root, rows, chg, want = cur[i]
                assert eng.score_batch([t], [Topology(left, right, root, [])]).tolist() == [want], (rnd, i)
                assert np.array_equal(eng.changes(t)[n:], chg[n:]), (rnd, i)
                for v in (n, n + 71, 2 * n - 4):
                    assert np.array_equal(eng.download_row(t, v), rows[v]), (rnd, i, v)
            if rnd == 3:  # a classic commit of tree 0 on top of flipped banks
                left, right, root, rows, chg, _ = cur[0]
                l2, r2, dirty = _nni_of(left, right, root, n, n + 40, True)
                rr, cc = rows.copy(), chg.copy()
                rr[dirty, 0] = 0
                want, rr, cc = oracles.oracle_getplen(enc, l2, r2, root, rr, cc)
                assert eng.commit(trees[0], Topology(l2, r2, root, dirty)) == want
                cur[0] = [l2, r2, root, rr, cc, want]
            if rnd == 7:  # the neighbourhood and a copy read the rows where they are
                left, right, root, rows, chg, want = cur[1]
                nn = eng.nni_neighbourhood(trees[1], scratch, Topology(left, right, root))
                l2, r2, _ = _nni_of(left, right, root, n, n + 9, False)
                assert nn[2 * 9 + 0] == oracles.oracle_getplen(enc, l2, r2, root)[0]
                spare = eng.tree_create()
                eng.tree_copy(spare, trees[1])
                assert eng.score_batch([spare], [Topology(left, right, root, [])]).tolist() == [want]
                assert np.array_equal(eng.download_row(spare, n + 71), rows[n + 71])
                eng.tree_destroy(spare)
    finally:
        eng.close()


def test_keep_and_accept_errors_are_loud(golden_dir):
    """misuse of the score-and-keep calls fails with a message, and the tree is as it was"""
    g = load(golden_dir, "example.npz")
    enc = g["enc"]
    n = enc.shape[0]
    eng = Engine(0)
    try:
        eng.upload(enc, max_trees=2)
        t = eng.tree_create()
        left, right, root = g["random_left"], g["random_right"], int(g["random_root"])
        base = Topology(left, right, root)
        cand = Topology(g["nni_left"], g["nni_right"], int(g["nni_root"]), np.nonzero(g["nni_dirty"])[0])
        with pytest.raises(GpuError, match="never been evaluated"):
            eng.score_keep([t], [cand], [1])
        assert eng.commit(t, base) == 11808
        with pytest.raises(GpuError, match="two kept candidates"):
            eng.score_keep([t, t], [cand, cand], [1, 1])
        lens, ticket = eng.score_keep([t, t], [cand, cand], [0, 1])
        assert lens.tolist() == [11834, 11834]
        with pytest.raises(GpuError, match="was not kept"):
            eng.accept_kept(ticket, [0])
        with pytest.raises(GpuError, match="was not kept"):
            eng.accept_kept(ticket, [7])
        plain = eng.score_batch_submit([t], [cand])
        assert eng.batch_wait(plain).tolist() == [11834]
        with pytest.raises(GpuError, match="does not name a batch"):
            eng.accept_kept(plain[0], [0])
        assert eng.score_batch([t], [Topology(left, right, root, [])]).tolist() == [11808]  # nothing has changed
        eng.accept_kept(ticket, [1])
        with pytest.raises(GpuError, match="accepted already"):
            eng.accept_kept(ticket, [1])
        assert eng.score_batch([t], [Topology(g["nni_left"], g["nni_right"], int(g["nni_root"]), [])]).tolist() == [11834]
        want, rows, chg = oracles.oracle_getplen(enc, g["nni_left"], g["nni_right"], int(g["nni_root"]))
        assert np.array_equal(eng.changes(t)[n:], chg[n:])
    finally:
        eng.close()


def test_trace_of_a_store_launch(monkeypatch):
    """LVB_GPU_TRACE=1 (the measuring aid behind tools/trace_store.py): every work item of a whole-tree store launch
    leaves six increasing time stamps, and the result of the launch is the oracle's"""
    import ctypes as C

    from lvb_b200 import load_library
    monkeypatch.setenv("LVB_GPU_TRACE", "1")
    n, m = 400, 60_000
    enc = synthetic.pack_states(synthetic.states(n, m, seed=4, p=0.1))
    left, right, root = synthetic.random_topology(n, seed=4)
    eng = Engine(0)
    try:
        eng.upload(enc, max_trees=1)
        t = eng.tree_create()
        assert eng.commit(t, Topology(left, right, root)) == oracles.oracle_getplen(enc, left, right, root)[0]
        lib = load_library()
        lib.lvb_gpu_trace_download.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.POINTER(C.c_int64), C.POINTER(C.c_int64)]
        ni, ncb = C.c_int64(0), C.c_int64(0)
        assert lib.lvb_gpu_trace_download(eng.ctx, None, 0, C.byref(ni), C.byref(ncb)) == 0
        assert ni.value > 0 and ncb.value > 0 and ni.value % ncb.value == 0
        tr = np.zeros((ni.value, 6), dtype=np.uint64)
        assert lib.lvb_gpu_trace_download(eng.ctx, tr.ctypes.data, ni.value, C.byref(ni), C.byref(ncb)) == 0
        assert (tr > 0).all()
        d = np.diff(tr.astype(np.int64), axis=1)
        assert (d[:, [0, 1, 3, 4]] >= 0).all()  # ticket <= decoded <= published; first operand <= last step <= books
    finally:
        eng.close()


def test_large_score_batch_two_columns_per_thread():
    """a score launch of thousands of CTAs goes through fitch_score2_kernel (two columns per thread): every NNI
    candidate of a 150-taxon tree on a ragged 200 003-site alignment (the last column block is partly empty, the
    second column of its threads wholly), against the oracle"""
    n, m = 150, 200_003
    enc = synthetic.pack_states(synthetic.states(n, m, seed=9, p=0.1))
    left, right, root = synthetic.random_topology(n, seed=9)
    eng = Engine(0)
    try:
        eng.upload(enc, max_trees=1)
        t = eng.tree_create()
        want, rows, chg = oracles.oracle_getplen(enc, left, right, root)
        assert eng.commit(t, Topology(left, right, root)) == want
        cands = nni_neighbourhood(left, right, root, n)
        assert len(cands) * ((enc.shape[1] // 2 + 511) // 512) >= 4 * 148 * 6  # enough CTAs for the two-column route
        got = eng.score_batch([t] * len(cands), cands)
        for k in range(len(cands)):
            if k % 3 == 0 or k > len(cands) - 4:
                assert got[k] == oracles.oracle_getplen(enc, cands[k].left, cands[k].right, root)[0], k
        # the same candidates in small batches (the one-column kernels) give the same lengths
        small = np.concatenate([eng.score_batch([t] * 6, cands[i:i + 6]) for i in range(0, 60, 6)])
        assert np.array_equal(small, got[:60])
    finally:
        eng.close()
