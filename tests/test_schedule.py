"""CPU: the host schedule compiler (lvb_b200/csrc/schedule.cpp, reached through the C-ABI's
lvb_gpu_compile_schedule) checked by interpreting its instruction stream in numpy
(tests/fitch_vm.py) against the reference's golden lengths / rows / per-node counts."""
import os

import numpy as np
import pytest

import fitch_vm
from lvb_b200 import GpuError, Topology, compile_schedule


def load(golden_dir, name):
    return np.load(os.path.join(golden_dir, name))


def fresh_rows(enc):
    n, nwords = enc.shape
    rows = np.zeros((2 * n - 3, nwords), dtype=np.uint64)
    rows[:n] = enc
    return rows


def test_full_pass_program_shape_and_result(golden_dir):
    g = load(golden_dir, "example.npz")
    enc = g["enc"]
    n = enc.shape[0]
    topo = Topology(g["random_left"], g["random_right"], int(g["random_root"]))
    instr, slots = compile_schedule(n, topo, store=True)
    assert instr.shape == (n - 3 + 2, 4)  # one per internal node + the two root-step applications
    assert 0 < slots <= 24
    nodes = instr[:, 2]
    assert sorted(nodes[nodes >= 0].tolist()) == list(range(n, 2 * n - 3))
    # every HBM operand is a leaf row: internal results travel in registers / stack slots only
    ops = instr[:, :2]
    assert ops.max() < n
    rows = fresh_rows(enc)
    total, per_node, rootc = fitch_vm.run_program(instr, rows, n, store=True)
    assert total == int(g["random_length"]) == 11808
    assert np.array_equal(rows, g["random_rows_after"])
    want = g["random_changes_after"]
    assert all(per_node[v] == want[v] for v in range(n, 2 * n - 3))
    assert rootc == 11808 - want[n:].sum()


@pytest.mark.parametrize("step,prev", [("nni", "random"), ("spr", "nni"), ("tbr", "spr")])
def test_incremental_program(golden_dir, step, prev):
    g = load(golden_dir, "example.npz")
    enc = g["enc"]
    n = enc.shape[0]
    rows = g[f"{prev}_rows_after"].copy()
    chg = g[f"{prev}_changes_after"]
    dirty = np.nonzero(g[f"{step}_dirty"])[0]
    assert dirty.min() >= n
    topo = Topology(g[f"{step}_left"], g[f"{step}_right"], int(g[f"{step}_root"]), dirty)
    # score: nothing stored, length = cached counts of clean nodes + recomputed + root step
    instr, slots = compile_schedule(n, topo, store=False)
    assert len(instr) == len(dirty) + 2
    before = rows.copy()
    total, per_node, rootc = fitch_vm.run_program(instr, rows, n, store=False)
    assert np.array_equal(rows, before)
    clean = np.setdiff1d(np.arange(n, 2 * n - 3), dirty)
    assert chg[clean].sum() + total == int(g[f"{step}_length"])
    # commit: same program with stores reproduces the reference's rows and counts
    instr, _ = compile_schedule(n, topo, store=True)
    total2, per_node2, _ = fitch_vm.run_program(instr, rows, n, store=True)
    assert total2 == total
    assert np.array_equal(rows, g[f"{step}_rows_after"])
    assert all(per_node2[v] == g[f"{step}_changes_after"][v] for v in dirty)


def test_proposal_series_scores(golden_dir):
    """300 reference proposals (nni/spr/tbr, accept-if-not-worse, re-root every 50): score each
    against the current state, commit the accepted ones; re-rooting recomputes only the nodes
    whose children changed"""
    g = load(golden_dir, "proposals.npz")
    enc = load(golden_dir, "example.npz")["enc"]
    n = enc.shape[0]
    B = 2 * n - 3
    rows = fresh_rows(enc)
    chg = np.zeros(B, dtype=np.int64)
    cur_l, cur_r, cur_root = g["start_left"].astype(np.int64), g["start_right"].astype(np.int64), int(g["start_root"])
    instr, _ = compile_schedule(n, Topology(cur_l, cur_r, cur_root), store=True)
    _, per_node, _ = fitch_vm.run_program(instr, rows, n, store=True)
    for v, c in per_node.items():
        chg[v] = c
    max_slots = 0
    for i in range(len(g["length"])):
        left, right, root = g["left"][i].astype(np.int64), g["right"][i].astype(np.int64), int(g["root"][i])
        ref_dirty = np.nonzero(g["dirty"][i])[0]
        # nodes whose children differ from the device's current tree (covers re-rooting, where
        # the reference dirties everything but only the path between the roots changes)
        changed = np.nonzero((left != cur_l) | (right != cur_r))[0]
        changed = changed[changed >= n]
        if len(ref_dirty) == n - 3:
            # all-dirty after a re-root: the candidate differs from the current tree along the
            # re-root path and the proposal's own paths; close the changed set towards the root
            parent = np.full(B, -1, dtype=np.int64)
            for v in range(n, B):
                parent[left[v]] = v
                parent[right[v]] = v
            parent[left[root]] = root
            parent[right[root]] = root
            dirty = set()
            for v in changed.tolist():
                while v >= n and v not in dirty:
                    dirty.add(v)
                    v = parent[v]
            dirty = np.array(sorted(dirty), dtype=np.int64)
        else:
            dirty = ref_dirty
            assert set(changed.tolist()) <= set(dirty.tolist())
        topo = Topology(left, right, root, dirty)
        instr, slots = compile_schedule(n, topo, store=False)
        max_slots = max(max_slots, slots)
        total, _, _ = fitch_vm.run_program(instr, rows, n, store=False)
        clean = np.setdiff1d(np.arange(n, B), dirty)
        assert chg[clean].sum() + total == int(g["length"][i]), i
        if g["accepted"][i]:
            instr, _ = compile_schedule(n, topo, store=True)
            _, per_node, _ = fitch_vm.run_program(instr, rows, n, store=True)
            for v, c in per_node.items():
                chg[v] = c
            cur_l, cur_r, cur_root = left, right, root
    assert max_slots <= 2  # SPR/TBR: two dirty paths meet once or twice


def test_slots_follow_tree_shape():
    # caterpillar: every internal node has a leaf child -> no stack at all
    n = 64
    B = 2 * n - 3
    left = np.full(B, -1, dtype=np.int64)
    right = np.full(B, -1, dtype=np.int64)
    # root 0; internal nodes n..B-1 chained; leaves 1..n-1
    left[0], right[0] = n, 1
    leaf = 2
    for v in range(n, B):
        if v + 1 < B:
            left[v], right[v] = v + 1, leaf
            leaf += 1
        else:
            left[v], right[v] = leaf, leaf + 1
    instr, slots = compile_schedule(n, Topology(left, right, 0), store=True)
    assert slots == 0 and len(instr) == n - 1
    # balanced tree over 64 leaves + the root leaf: log2 depth of stack
    n = 65
    B = 2 * n - 3
    left = np.full(B, -1, dtype=np.int64)
    right = np.full(B, -1, dtype=np.int64)
    level = list(range(1, n))
    nxt = n
    while len(level) > 2:
        new = []
        for k in range(0, len(level), 2):
            left[nxt], right[nxt] = level[k], level[k + 1]
            new.append(nxt)
            nxt += 1
        level = new
    left[0], right[0] = level
    assert nxt == B
    instr, slots = compile_schedule(n, Topology(left, right, 0), store=True)
    assert slots == 5 and len(instr) == n - 1  # 32 cherries -> ... -> 2 subtrees: depth 5


def test_malformed_topologies_are_rejected(golden_dir):
    g = load(golden_dir, "example.npz")
    n = g["enc"].shape[0]
    left, right = g["random_left"].copy(), g["random_right"].copy()
    with pytest.raises(GpuError):
        compile_schedule(n, Topology(left, right, n + 1))  # root must be a leaf
    with pytest.raises(GpuError):
        compile_schedule(n, Topology(left, right, 0, [3]))  # a leaf cannot be dirty
    bad = left.copy()
    bad[n + 5] = n + 5  # self loop
    with pytest.raises(GpuError):
        compile_schedule(n, Topology(bad, right, 0))
    bad = left.copy()
    v = n + 7
    bad[v] = right[v]  # both children equal
    with pytest.raises(GpuError):
        compile_schedule(n, Topology(bad, right, 0))
    # a node reachable from two parents
    bad = left.copy()
    kids = [int(c) for c in left[n:] if c >= n]
    bad[n + 1 if left[n + 1] < n else n + 2] = kids[0]
    with pytest.raises(GpuError):
        compile_schedule(n, Topology(bad, right, 0))


def test_empty_dirty_set_is_just_the_root_step(golden_dir):
    g = load(golden_dir, "example.npz")
    n = g["enc"].shape[0]
    rows = g["random_rows_after"].copy()
    instr, slots = compile_schedule(n, Topology(g["random_left"], g["random_right"], 0, []), store=False)
    assert len(instr) == 2 and slots == 0
    total, _, rootc = fitch_vm.run_program(instr, rows, n, store=False)
    assert total == rootc == 11808 - g["random_changes_after"][n:].sum()


@pytest.mark.parametrize("piece_size", [1, 3, 8, 16, 40, 97, 1000])
def test_pieces_cover_the_program_and_respect_levels(golden_dir, piece_size):
    """a store program cut into independent subtrees: every node exactly once, pieces of a
    level only read rows stored by earlier levels, result identical to the uncut program"""
    from lvb_b200 import compile_pieces

    g = load(golden_dir, "example.npz")
    enc = g["enc"]
    n = enc.shape[0]
    topo = Topology(g["tbr_left"], g["tbr_right"], int(g["tbr_root"]))
    instr, pieces, slots = compile_pieces(n, topo, piece_size)
    assert len(instr) == n - 1
    assert pieces[:, 1].sum() == len(instr) and (pieces[:, 1] > 0).all()
    nodes = instr[:, 2]
    assert sorted(nodes[nodes >= 0].tolist()) == list(range(n, 2 * n - 3))
    if piece_size >= n:
        assert len(pieces) == 1
    else:
        # level 0: forests of maximal subtrees of at most piece_size nodes, packed largest-first into bins
        # that are closed at 5/4 of their capacity (piece_size, smaller towards the end, never under 4)
        lvl0 = pieces[pieces[:, 2] == 0, 1]
        assert (lvl0 <= max(5, 5 * piece_size // 4)).all()
        assert pieces[-1, 2] == pieces[:, 2].max() and (np.diff(pieces[:, 2]) >= 0).all()
        assert set(pieces[:, 2].tolist()) == set(range(pieces[-1, 2] + 1))  # levels are contiguous
    # run level by level; within a level, pieces may only read leaves and rows of lower levels
    rows = fresh_rows(enc)
    stored_level = {}
    total = 0
    for lvl in sorted(set(pieces[:, 2].tolist())):
        wrote = []
        for begin, cnt, plvl in pieces.tolist():
            if plvl != lvl:
                continue
            part = instr[begin:begin + cnt]
            for a, b, node, ctl in part.tolist():
                for op in (a, b):
                    if op >= n:
                        assert stored_level[op] < lvl, (op, lvl)
                if node >= 0:
                    wrote.append(node)
            t, _, _ = fitch_vm.run_program(part, rows, n, store=True)
            total += t
        for v in wrote:
            stored_level[v] = lvl
    assert total == int(g["tbr_length"])
    assert np.array_equal(rows, g["tbr_rows_after"])


def _nni_candidates(left, right, root, n):
    """mutate_deterministic (TreeOperations.c:105-158) for every internal node and both sides,
    in the order of deterministic_hillclimb's loops (Solve.c:100-103); plain test-input code"""
    B = 2 * n - 3
    parent = np.full(B, -1, dtype=np.int64)
    for v in list(range(n, B)) + [root]:
        parent[left[v]] = v
        parent[right[v]] = v
    for u in range(n, B):
        v = int(parent[u])
        a, b = int(left[u]), int(right[u])
        c = int(right[v]) if left[v] == u else int(left[v])
        for take_left in (False, True):
            l2, r2 = left.copy(), right.copy()
            moved = a if take_left else b
            if l2[v] == u:
                r2[v] = moved
            else:
                l2[v] = moved
            l2[u], r2[u] = (b, c) if take_left else (a, c)
            yield u, v, a, b, c, l2, r2


@pytest.mark.parametrize("n,m,shape", [(5, 40, "random"), (6, 50, "random"), (37, 300, "random"), (120, 200, "random"),
                                       (40, 100, "caterpillar"), (33, 100, "balanced"), (40, 100, "balanced")])
def test_up_pass_program_and_four_row_nni_lengths(n, m, shape):
    """The up-pass program (pre-order, previous result / stack slots) run by the numpy VM gives,
    with the four-row evaluation on the rearranged branch, exactly the length a full oracle pass
    returns for every mutate_deterministic candidate — the identity lvb_gpu_nni_neighbourhood
    rests on (Fitch lengths do not depend on the root)."""
    import oracles
    from lvb_b200 import compile_up_pass, synthetic

    rng = np.random.default_rng(n * 1000 + m)
    st = rng.integers(1, 16, size=(n, m), dtype=np.uint8) if shape != "random" else synthetic.states(n, m, seed=n, p=0.15)
    enc = synthetic.pack_states(st)
    left, right, root = synthetic.random_topology(n, seed=n + 1) if shape == "random" else synthetic.shaped_topology(n, shape)
    base, rows, _ = oracles.oracle_getplen(enc, left, right, root)
    instr, slots, pieces = compile_up_pass(n, Topology(left, right, root))
    B = 2 * n - 3
    assert instr.shape == (n - 3, 4) and pieces.tolist() == [[0, n - 3, 0]]
    assert sorted(instr[:, 2].tolist()) == list(range(n, B))      # every internal node once
    assert np.all(instr[:, 3] & fitch_vm.CTL_STORE)
    assert slots <= max(1, int(np.ceil(np.log2(n))))
    # no instruction reads a row the program stores: internal operands are rows of the evaluated tree
    up = rows.copy()
    fitch_vm.run_program(instr, _ReadLowWriteUp(rows, up, B, n), n, store=True)
    # cut into a top piece and independent subtrees: same up sets, level 1 after level 0, any order inside
    for piece_size in (2, 5, 16):
        pi, _, pcs = compile_up_pass(n, Topology(left, right, root), piece_size)
        assert pcs[0].tolist()[2] == 0 and all(p[2] == 1 for p in pcs[1:].tolist())
        assert sum(p[1] for p in pcs.tolist()) == len(pi) == (n - 3) + (len(pcs) - 1)   # one loader per cut piece
        stored = pi[pi[:, 2] >= 0]
        assert sorted(stored[:, 2].tolist()) == list(range(n, B))
        up2 = np.zeros_like(rows)
        view = _ReadLowWriteUp(rows, up2, B, n)
        for b, k, _lvl in [pcs[0].tolist()] + pcs[1:][::-1].tolist():
            fitch_vm.run_program(pi[b:b + k], view, n, store=True)
        assert np.array_equal(up2[n:], up[n:]), piece_size
    for u, v, a, b, c, l2, r2 in _nni_candidates(left, right, root, n):
        A, Bq, Cq = rows[a], rows[b], rows[c]
        D = rows[root] if v == root else up[v]

        def join(p, q, r, s):
            pq, c1 = fitch_vm.fitch_rows(p, q)
            rs, c2 = fitch_vm.fitch_rows(r, s)
            return c1 + c2 + fitch_vm.fitch_rows(pq, rs)[1]

        cur = join(A, Bq, Cq, D)
        got = base - cur + (join(Bq, Cq, A, D) if l2[u] == b else join(A, Cq, Bq, D))
        assert got == oracles.oracle_getplen(enc, l2, r2, root)[0], (u, l2[u] == b)


class _ReadLowWriteUp:
    """rows[] as the up-pass sees them: operands come from the evaluated tree, stores go to the scratch tree"""

    def __init__(self, low, up, B, n):
        self.low, self.up, self.B, self.n = low, up, B, n

    def __getitem__(self, v):
        # B + (v - n): the row of v in the tree the program stores to (fitch_isa.h)
        return self.up[v - self.B + self.n] if v >= self.B else self.low[v]

    def __setitem__(self, v, val):
        self.up[v] = val
