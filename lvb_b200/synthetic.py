"""Synthetic DNA alignments of the BASELINE.json shapes (data generation only, numpy).

Sequences evolve down a random bifurcating tree as in SURVEY.md appendix A (root sequence
uniform over ACGT, each site of each child substituted with probability p), so that parsimony
lengths look like real data; rows are shuffled.  `packed()` returns the 4-bit-per-site words
in LVB's layout (site k of a word at bits 4k..4k+3, A=1 C=2 G=4 T=8, tail padded with 15 —
/root/reference/src/DataOperations.c:164-249) without going through text.
"""
from __future__ import annotations

import numpy as np


def states(n: int, m: int, seed: int = 1, p: float = 0.02) -> np.ndarray:
    """n x m uint8 matrix of states 0..3 (A,C,G,T)."""
    rng = np.random.default_rng(seed)
    seqs = [rng.integers(0, 4, size=m, dtype=np.uint8)]
    while len(seqs) < n:
        parent = seqs.pop(int(rng.integers(0, len(seqs))))
        for _ in range(2):
            child = parent.copy()
            idx = np.nonzero(rng.random(m, dtype=np.float32) < p)[0]
            child[idx] = (child[idx] + rng.integers(1, 4, size=idx.size, dtype=np.uint8)) % 4
            seqs.append(child)
    order = rng.permutation(n)
    return np.stack([seqs[i] for i in order])


def pack_states(st: np.ndarray) -> np.ndarray:
    """uint8 states 0..3 -> LVB packed words, n x ceil(m/16) uint64."""
    n, m = st.shape
    nwords = (m + 15) // 16
    out = np.empty((n, nwords), dtype=np.uint64)
    shifts = (np.arange(16, dtype=np.uint64) * np.uint64(4))[None, :]
    for i in range(n):
        nib = np.full(nwords * 16, 15, dtype=np.uint64)
        nib[:m] = np.uint64(1) << st[i].astype(np.uint64)
        out[i] = np.bitwise_or.reduce(nib.reshape(nwords, 16) << shifts, axis=1)
    return out


def text_rows(st: np.ndarray):
    lut = np.frombuffer(b"ACGT", dtype=np.uint8)
    return [bytes(lut[r]) for r in st]


def packed(n: int, m: int, seed: int = 1, p: float = 0.02) -> np.ndarray:
    return pack_states(states(n, m, seed, p))


def random_packed(n: int, nwords: int, seed: int = 1) -> np.ndarray:
    """Fast unstructured alternative for very large shapes: every site an independent uniform
    base (one-hot nibble).  Same memory traffic and instruction mix as real data (the kernels
    have no data-dependent control flow); lengths are simply larger."""
    rng = np.random.default_rng(seed)
    # one byte = two sites: low nibble 1<<a, high nibble 1<<b
    lut = np.array([(1 << (k & 3)) | (16 << (k >> 2)) for k in range(16)], dtype=np.uint8)
    out = np.empty((n, nwords), dtype=np.uint64)
    for i in range(n):
        out[i] = lut[rng.integers(0, 16, size=nwords * 8, dtype=np.uint8)].view(np.uint64)
    return out


def random_topology(n: int, seed: int = 1):
    """A random rooted-at-leaf-0 binary tree in LVB numbering (leaves 0..n-1, internal
    n..2n-4), grown by splitting random leaves.  Returns (left, right, root) int64 arrays.
    For tests and benches that need some tree; not LVB's PullRandomTree stream."""
    rng = np.random.default_rng(seed)
    B = 2 * n - 3
    left = np.full(B, -1, dtype=np.int64)
    right = np.full(B, -1, dtype=np.int64)
    # start: root 0 with children 1, 2 (all leaves); grow by replacing a random non-root leaf
    # slot with a new internal node
    leaves = [1, 2]
    parent_slot = {1: (0, "l"), 2: (0, "r")}
    left[0], right[0] = 1, 2
    next_leaf, next_int = 3, n
    while next_leaf < n:
        k = int(rng.integers(0, len(leaves)))
        v = leaves[k]
        par, side = parent_slot[v]
        node = next_int
        next_int += 1
        if side == "l":
            left[par] = node
        else:
            right[par] = node
        w = next_leaf
        next_leaf += 1
        left[node], right[node] = v, w
        parent_slot[node] = (par, side)
        parent_slot[v] = (node, "l")
        parent_slot[w] = (node, "r")
        leaves.append(w)
    assert next_int == B
    return left, right, 0


def shaped_topology(n: int, shape: str):
    """extreme tree shapes, root = leaf 0: "caterpillar" (a chain of n-3 internal nodes) or
    "balanced" (leaves paired level by level)"""
    B = 2 * n - 3
    left, right = np.full(B, -1, dtype=np.int64), np.full(B, -1, dtype=np.int64)
    if shape == "caterpillar":
        left[0], right[0] = n, 1
        leaf = 2
        for v in range(n, B):
            if v + 1 < B:
                left[v], right[v] = v + 1, leaf
                leaf += 1
            else:
                left[v], right[v] = leaf, leaf + 1
        return left, right, 0
    if shape != "balanced":
        raise ValueError(shape)
    level, nxt = list(range(1, n)), n
    while len(level) > 2:
        new = []
        for k in range(0, len(level) - 1, 2):
            left[nxt], right[nxt] = level[k], level[k + 1]
            new.append(nxt)
            nxt += 1
        if len(level) % 2:
            new.append(level[-1])
        level = new
    left[0], right[0] = level
    assert nxt == B
    return left, right, 0


def survey_alignment(n: int, m: int, seed: int = 1, p: float = 0.05) -> np.ndarray:
    """The exact generator of SURVEY.md appendix A (float64 draws), n x m uint8 states 0..3 in
    file order.  `survey_alignment(100, 10000, 1, 0.05)` written by `write_phylip` is the file
    whose reference run is recorded there (md5 a11fa9393aaf6c6b8150b4491e8a4d90)."""
    rng = np.random.default_rng(seed)
    seqs = [rng.integers(0, 4, size=m, dtype=np.uint8)]
    while len(seqs) < n:
        parent = seqs.pop(int(rng.integers(0, len(seqs))))
        for _ in range(2):
            child = parent.copy()
            mask = rng.random(m) < p
            child[mask] = (child[mask] + rng.integers(1, 4, size=int(mask.sum()), dtype=np.uint8)) % 4
            seqs.append(child)
    order = rng.permutation(n)
    return np.stack([seqs[i] for i in order])


def write_phylip(path: str, st: np.ndarray) -> None:
    """PHYLIP sequential, one line per taxon, names T<k> left-justified to 10 characters."""
    n, m = st.shape
    lut = np.frombuffer(b"ACGT", dtype=np.uint8)
    with open(path, "wb") as f:
        f.write(f" {n} {m}\n".encode())
        for k in range(n):
            f.write(f"T{k:<9d}".encode() + bytes(lut[st[k]]) + b"\n")
