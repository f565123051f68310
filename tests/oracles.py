"""Test infrastructure: ctypes access to the oracle (oracle/liboracle.so, our plain-C
restatement) and, when built, to the real reference (oracle/_ref/libLVB_ref.so).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline/reference arm import this.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_SO = os.path.join(ROOT, "oracle", "liboracle.so")
REF_SO = os.path.join(ROOT, "oracle", "_ref", "libLVB_ref.so")
REF_BIN = os.path.join(ROOT, "oracle", "_ref", "lvb_ref")

_vp, _l, _lp = C.c_void_p, C.c_long, C.POINTER(C.c_long)


def build_oracle() -> None:
    subprocess.run(["make", "-s", "-C", os.path.join(ROOT, "oracle"), "all"], check=True,
                   stdout=subprocess.DEVNULL)


_oracle = None


def oracle() -> C.CDLL:
    global _oracle
    if _oracle is None:
        if not os.path.exists(ORACLE_SO):
            build_oracle()
        L = C.CDLL(ORACLE_SO)
        L.oracle_words_per_row.restype = _l
        L.oracle_words_per_row.argtypes = [_l]
        L.oracle_dna_to_binary.argtypes = [_l, _l, C.POINTER(C.c_char_p), _vp]
        L.oracle_variable_columns.restype = _l
        L.oracle_variable_columns.argtypes = [_l, _l, C.POINTER(C.c_char_p), _vp]
        L.oracle_min_tree_length.restype = _l
        L.oracle_min_tree_length.argtypes = [_l, _l, C.POINTER(C.c_char_p)]
        L.oracle_fitch_word.argtypes = [C.c_uint64, C.c_uint64, C.POINTER(C.c_uint64)]
        L.oracle_getplen.restype = _l
        L.oracle_getplen.argtypes = [_l, _l, _vp, _vp, _l, _vp, _vp]
        L.oracle_getplen_sliced.restype = _l
        L.oracle_getplen_sliced.argtypes = [_l, _l, _vp, _vp, _l, _vp, _vp, C.c_int]
        L.oracle_dirty_all.argtypes = [_l, _l, _vp]
        _oracle = L
    return _oracle


def have_ref() -> bool:
    return os.path.exists(REF_SO)


_ref = None


def ref() -> C.CDLL:
    global _ref
    if _ref is None:
        L = C.CDLL(REF_SO)
        L.ref_msa_load.restype = _vp
        L.ref_msa_load.argtypes = [C.c_char_p, C.c_int, C.c_int]
        L.ref_msa_from_rows.restype = _vp
        L.ref_msa_from_rows.argtypes = [_l, _l, C.POINTER(C.c_char_p), C.c_int]
        L.ref_msa_from_packed.restype = _vp
        L.ref_msa_from_packed.argtypes = [_l, _l, _vp, C.c_int]
        L.ref_msa_free.argtypes = [_vp]
        L.ref_msa_dims.argtypes = [_vp, _lp]
        L.ref_msa_set_threads.argtypes = [_vp, C.c_int]
        L.ref_msa_row.argtypes = [_vp, _l, _vp]
        L.ref_dna_to_binary.argtypes = [_vp, _vp]
        L.ref_min_tree_length.restype = _l
        L.ref_min_tree_length.argtypes = [_vp]
        L.ref_rinit.argtypes = [C.c_int]
        L.ref_uni.restype = C.c_double
        L.ref_randpint.restype = _l
        L.ref_randpint.argtypes = [_l]
        L.ref_tree_alloc.restype = _vp
        L.ref_tree_alloc.argtypes = [_vp]
        L.ref_tree_free.argtypes = [_vp]
        L.ref_pull_random_tree.argtypes = [_vp, _vp]
        L.ref_ss_init.argtypes = [_vp, _vp]
        L.ref_getplen.restype = _l
        L.ref_getplen.argtypes = [_vp, _vp, _l]
        L.ref_treecopy.argtypes = [_vp, _vp, _vp]
        for f in (L.ref_mutate_nni, L.ref_mutate_spr, L.ref_mutate_tbr):
            f.argtypes = [_vp, _vp, _vp, _l]
        L.ref_mutate_deterministic.argtypes = [_vp, _vp, _vp, _l, _l, C.c_int]
        L.ref_arbreroot.restype = _l
        L.ref_arbreroot.argtypes = [_vp, _vp, _l]
        L.ref_lvb_reroot.restype = _l
        L.ref_lvb_reroot.argtypes = [_vp, _vp, _l, _l]
        L.ref_tree_dirty_all.argtypes = [_vp, _vp]
        L.ref_tree_get.argtypes = [_vp, _vp, _vp, _vp, _vp, _vp, _vp]
        L.ref_tree_set_topology.argtypes = [_vp, _vp, _vp, _vp, _vp]
        L.ref_tree_rows.argtypes = [_vp, _vp, _vp]
        L.ref_starting_temperature.restype = C.c_double
        L.ref_starting_temperature.argtypes = [_vp, _vp, _l]
        L.ref_deterministic_hillclimb.argtypes = [_vp, _vp, _l, C.c_char_p, _lp]
        L.ref_time_full_passes.restype = C.c_double
        L.ref_time_full_passes.argtypes = [_vp, _vp, _l, C.c_int, _lp]
        L.ref_time_proposals.restype = C.c_double
        L.ref_time_proposals.argtypes = [_vp, _vp, _vp, _l, C.c_int, C.c_int, _vp, C.POINTER(C.c_double), _lp]
        _ref = L
    return _ref


class RefMSA:
    """An alignment held by the reference (after matchange, with DNAToBinary done)."""

    def __init__(self, handle):
        self.L = ref()
        self.h = handle
        d = (C.c_long * 10)()
        self.L.ref_msa_dims(self.h, d)
        (self.n, self.m, self.original_m, self.nwords, self.bytes, self.B, self.min_len,
         self.n_threads, self.slice, self.tree_bytes) = list(d)

    @classmethod
    def load(cls, path: str, fmt: int = 0, threads: int = 1) -> "RefMSA":
        return cls(ref().ref_msa_load(path.encode(), fmt, threads))

    @classmethod
    def from_rows(cls, rows, threads: int = 1) -> "RefMSA":
        rows = [r if isinstance(r, bytes) else bytes(r) for r in rows]
        arr = (C.c_char_p * len(rows))(*rows)
        return cls(ref().ref_msa_from_rows(len(rows), len(rows[0]), arr, threads))

    @classmethod
    def from_packed(cls, enc: np.ndarray, m: int, threads: int = 1) -> "RefMSA":
        """an alignment that is already packed (no text, no matchange): timing at large shapes"""
        enc = np.ascontiguousarray(enc, dtype=np.uint64)
        assert enc.shape[1] == (m + 15) // 16
        return cls(ref().ref_msa_from_packed(enc.shape[0], m, enc.ctypes.data, threads))

    def set_threads(self, threads: int) -> None:
        self.L.ref_msa_set_threads(self.h, threads)
        d = (C.c_long * 10)()
        self.L.ref_msa_dims(self.h, d)
        self.n_threads, self.slice = d[7], d[8]

    def rows_text(self):
        out = []
        for i in range(self.n):
            b = C.create_string_buffer(self.m)
            self.L.ref_msa_row(self.h, i, b)
            out.append(b.raw)
        return out

    def enc(self) -> np.ndarray:
        out = np.zeros((self.n, self.nwords), dtype=np.uint64)
        self.L.ref_dna_to_binary(self.h, out.ctypes.data)
        return out

    def tree(self) -> "RefTree":
        return RefTree(self)


class RefTree:
    def __init__(self, msa: RefMSA):
        self.msa, self.L = msa, msa.L
        self.t = self.L.ref_tree_alloc(msa.h)
        self.root = 0

    def random(self) -> "RefTree":
        self.L.ref_pull_random_tree(self.msa.h, self.t)
        self.L.ref_ss_init(self.msa.h, self.t)
        self.root = 0
        return self

    def getplen(self) -> int:
        return self.L.ref_getplen(self.msa.h, self.t, self.root)

    def mutate(self, kind: str, src: "RefTree") -> "RefTree":
        fn = {"nni": self.L.ref_mutate_nni, "spr": self.L.ref_mutate_spr, "tbr": self.L.ref_mutate_tbr}[kind]
        fn(self.msa.h, self.t, src.t, src.root)
        self.root = src.root
        return self

    def mutate_deterministic(self, src: "RefTree", p: int, left: bool) -> "RefTree":
        self.L.ref_mutate_deterministic(self.msa.h, self.t, src.t, src.root, p, int(left))
        self.root = src.root
        return self

    def copy_from(self, src: "RefTree") -> "RefTree":
        self.L.ref_treecopy(self.msa.h, self.t, src.t)
        self.root = src.root
        return self

    def arbreroot(self) -> int:
        self.root = self.L.ref_arbreroot(self.msa.h, self.t, self.root)
        return self.root

    def reroot(self, newroot: int) -> None:
        self.L.ref_lvb_reroot(self.msa.h, self.t, self.root, newroot)
        self.root = newroot

    def dirty_all(self) -> None:
        self.L.ref_tree_dirty_all(self.msa.h, self.t)

    def get(self):
        B = self.msa.B
        parent, left, right, changes = (np.zeros(B, dtype=np.int64) for _ in range(4))
        dirty = np.zeros(B, dtype=np.int32)
        self.L.ref_tree_get(self.msa.h, self.t, parent.ctypes.data, left.ctypes.data, right.ctypes.data,
                            changes.ctypes.data, dirty.ctypes.data)
        return parent, left, right, changes, dirty

    def set_topology(self, parent, left, right, root: int) -> None:
        parent, left, right = (np.ascontiguousarray(a, dtype=np.int64) for a in (parent, left, right))
        self.L.ref_ss_init(self.msa.h, self.t)
        self.L.ref_tree_set_topology(self.msa.h, self.t, parent.ctypes.data, left.ctypes.data, right.ctypes.data)
        self.root = root

    def rows(self) -> np.ndarray:
        out = np.zeros((self.msa.B, self.msa.nwords), dtype=np.uint64)
        self.L.ref_tree_rows(self.msa.h, self.t, out.ctypes.data)
        return out


# ---- plain-C oracle helpers ----
def oracle_pack(rows) -> np.ndarray:
    L = oracle()
    rows = [r if isinstance(r, bytes) else bytes(r) for r in rows]
    n, m = len(rows), len(rows[0])
    nwords = L.oracle_words_per_row(m)
    out = np.zeros((n, nwords), dtype=np.uint64)
    arr = (C.c_char_p * n)(*rows)
    rc = L.oracle_dna_to_binary(n, m, arr, out.ctypes.data)
    if rc != 0:
        raise ValueError("bad base symbol in data MSA")
    return out


def oracle_getplen(enc: np.ndarray, left, right, root: int, rows: np.ndarray | None = None,
                   changes: np.ndarray | None = None, slices: int = 0):
    """Run the oracle's getplen.  rows: B x nwords state (dirty = first word 0); if None a
    fresh tree is made (leaf rows from enc, every internal node dirty).  Returns
    (length, rows, changes)."""
    L = oracle()
    n, nwords = enc.shape
    B = 2 * n - 3
    if rows is None:
        rows = np.zeros((B, nwords), dtype=np.uint64)
        rows[:n] = enc
    if changes is None:
        changes = np.zeros(B, dtype=np.int64)
    left = np.ascontiguousarray(left, dtype=np.int64)
    right = np.ascontiguousarray(right, dtype=np.int64)
    if slices > 0:
        length = L.oracle_getplen_sliced(n, nwords, left.ctypes.data, right.ctypes.data, root, rows.ctypes.data,
                                         changes.ctypes.data, slices)
    else:
        length = L.oracle_getplen(n, nwords, left.ctypes.data, right.ctypes.data, root, rows.ctypes.data,
                                  changes.ctypes.data)
    return int(length), rows, changes
