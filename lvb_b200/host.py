"""ctypes binding of lvb_b200/lib/liblvbhost.so — the C++ host layer (lvb_b200/host/) that mirrors
LVB's own search, tree and alignment functions above the C-ABI of include/lvb_gpu.h.

Plumbing only.  Everything that returns a tree length goes to the GPU library; without it (or
without a GPU) those calls raise.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Optional, Sequence

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
HOST_LIB_PATH = os.path.join(_HERE, "lib", "liblvbhost.so")
CLI_PATH = os.path.join(_HERE, "bin", "lvb_gpu")

_lib = None
_vp, _l, _lp = C.c_void_p, C.c_long, C.POINTER(C.c_long)


class HostError(RuntimeError):
    """crash() of the host layer (the binary prints 'FATAL ERROR: ...' and exits instead)."""


def load_host_library() -> C.CDLL:
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(HOST_LIB_PATH):
        raise ImportError(f"{HOST_LIB_PATH} not found: build it with `make -C lvb_b200/host`")
    L = C.CDLL(HOST_LIB_PATH)
    L.lvbh_last_error.restype = C.c_char_p
    L.lvbh_rng_stream.argtypes = [C.c_int, _l, _vp, _l, _vp]
    L.lvbh_msa_from_rows.restype = _vp
    L.lvbh_msa_from_rows.argtypes = [_l, _l, C.POINTER(C.c_char_p)]
    L.lvbh_msa_load.restype = _vp
    L.lvbh_msa_load.argtypes = [C.c_char_p, C.c_int]
    L.lvbh_msa_free.argtypes = [_vp]
    L.lvbh_msa_dims.argtypes = [_vp, _lp]
    L.lvbh_msa_row.argtypes = [_vp, _l, _vp]
    L.lvbh_msa_title.argtypes = [_vp, _l, _vp, _l]
    L.lvbh_dna_to_binary.argtypes = [_vp, _vp]
    L.lvbh_rinit.argtypes = [_vp, C.c_int]
    L.lvbh_uni.restype = C.c_double
    L.lvbh_uni.argtypes = [_vp]
    L.lvbh_tree_alloc.restype = _vp
    L.lvbh_tree_alloc.argtypes = [_vp]
    L.lvbh_tree_free.argtypes = [_vp]
    L.lvbh_pull_random_tree.argtypes = [_vp, _vp]
    L.lvbh_tree_mark_clean.argtypes = [_vp, _vp]
    L.lvbh_mutate.argtypes = [_vp, C.c_int, _vp, _vp, _l]
    L.lvbh_mutate_deterministic.argtypes = [_vp, _vp, _vp, _l, _l, C.c_int]
    L.lvbh_arbreroot.restype = _l
    L.lvbh_arbreroot.argtypes = [_vp, _vp, _l, _vp, _lp]
    L.lvbh_reroot.argtypes = [_vp, _vp, _l, _l]
    L.lvbh_treecopy.argtypes = [_vp, _vp, _vp]
    L.lvbh_tree_get.argtypes = [_vp, _vp, _vp, _vp, _vp, _vp]
    L.lvbh_tree_set.argtypes = [_vp, _vp, _vp, _vp, _vp]
    L.lvbh_same_topology.argtypes = [_vp, _vp, _l, _vp, _l]
    L.lvbh_treeprint.argtypes = [_vp, _vp, _l, C.c_char_p]
    L.lvbh_device_open.argtypes = [_vp, C.c_int, C.c_int]
    L.lvbh_getplen.restype = _l
    L.lvbh_getplen.argtypes = [_vp, _vp, _l]
    L.lvbh_hillclimb.argtypes = [_vp, C.c_int, _vp, _l, C.c_char_p, _lp]
    L.lvbh_search.argtypes = [_vp, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, _l, C.c_int, C.c_int, C.c_char_p,
                              C.c_char_p, _vp, _vp, C.POINTER(C.c_double)]
    _lib = L
    return L


def _check(rc: int) -> None:
    if rc != 0:
        raise HostError(load_host_library().lvbh_last_error().decode(errors="replace"))


def rng_stream(seed: int, count: int, upper: int):
    """`count` uni() values followed by `count` randpint(upper) values from rinit(seed)."""
    u = np.zeros(count, dtype=np.float64)
    r = np.zeros(count, dtype=np.int64)
    _check(load_host_library().lvbh_rng_stream(seed, count, u.ctypes.data, upper, r.ctypes.data))
    return u, r


class HostMSA:
    """An alignment after matchange (DataStructure of the host layer) plus one random stream."""

    def __init__(self, handle):
        self.L = load_host_library()
        if not handle:
            raise HostError(self.L.lvbh_last_error().decode(errors="replace"))
        self.h = handle
        d = (C.c_long * 7)()
        self.L.lvbh_msa_dims(self.h, d)
        self.n, self.m, self.original_m, self.nwords, self.bytes, self.B, self.min_len = list(d)

    @classmethod
    def from_rows(cls, rows: Sequence[bytes]) -> "HostMSA":
        rows = [r if isinstance(r, bytes) else bytes(r) for r in rows]
        arr = (C.c_char_p * len(rows))(*rows)
        return cls(load_host_library().lvbh_msa_from_rows(len(rows), len(rows[0]), arr))

    @classmethod
    def load(cls, path: str, fmt: int = 0) -> "HostMSA":
        return cls(load_host_library().lvbh_msa_load(path.encode(), fmt))

    def close(self) -> None:
        if self.h:
            self.L.lvbh_msa_free(self.h)
            self.h = None

    def rows_text(self):
        out = []
        for i in range(self.n):
            b = C.create_string_buffer(self.m)
            self.L.lvbh_msa_row(self.h, i, b)
            out.append(b.raw)
        return out

    def title(self, i: int) -> str:
        b = C.create_string_buffer(256)
        self.L.lvbh_msa_title(self.h, i, b, 256)
        return b.value.decode()

    def enc(self) -> np.ndarray:
        out = np.zeros((self.n, self.nwords), dtype=np.uint64)
        _check(self.L.lvbh_dna_to_binary(self.h, out.ctypes.data))
        return out

    def rinit(self, seed: int) -> None:
        self.L.lvbh_rinit(self.h, seed)

    def uni(self) -> float:
        return float(self.L.lvbh_uni(self.h))

    def tree(self) -> "HostTree":
        return HostTree(self)

    def search(self, seed: int, algorithm: int = 1, cooling: int = 0, device: int = 0, max_trees: int = 0,
               replicates: int = 1, speculation: int = 0, max_rearrangements: int = 0, host_threads: int = 1,
               groups: int = 0, tsv_path: Optional[str] = None, outtree_path: Optional[str] = None):
        """GetSoln for `replicates` chains; returns a list of dicts (one per chain)."""
        out = np.zeros(7 * replicates, dtype=np.int64)
        t0 = np.zeros(replicates, dtype=np.float64)
        secs = C.c_double(0)
        _check(self.L.lvbh_search(self.h, device, seed, algorithm, cooling, max_trees, replicates, speculation,
                                  max_rearrangements, host_threads, groups, tsv_path.encode() if tsv_path else None,
                                  outtree_path.encode() if outtree_path else None, out.ctypes.data, t0.ctypes.data,
                                  C.byref(secs)))
        keys = ("final_length", "rearrangements", "topologies", "initial_length", "trees_scored", "score_launches", "groups")
        res = []
        for r in range(replicates):
            d = {k: int(out[7 * r + i]) for i, k in enumerate(keys)}
            d["start_temperature"] = float(t0[r])
            d["seconds"] = float(secs.value)
            res.append(d)
        return res


class HostTree:
    def __init__(self, msa: HostMSA):
        self.msa, self.L = msa, msa.L
        self.t = self.L.lvbh_tree_alloc(msa.h)
        self.root = 0

    def __del__(self):
        try:
            self.L.lvbh_tree_free(self.t)
        except Exception:
            pass

    def random(self) -> "HostTree":
        _check(self.L.lvbh_pull_random_tree(self.msa.h, self.t))
        self.root = 0
        return self

    def mark_clean(self) -> "HostTree":
        self.L.lvbh_tree_mark_clean(self.msa.h, self.t)
        return self

    def mutate(self, kind: str, src: "HostTree") -> "HostTree":
        _check(self.L.lvbh_mutate(self.msa.h, {"nni": 0, "spr": 1, "tbr": 2}[kind], self.t, src.t, src.root))
        self.root = src.root
        return self

    def mutate_deterministic(self, src: "HostTree", p: int, left: bool) -> "HostTree":
        _check(self.L.lvbh_mutate_deterministic(self.msa.h, self.t, src.t, src.root, p, int(left)))
        self.root = src.root
        return self

    def copy_from(self, src: "HostTree") -> "HostTree":
        self.L.lvbh_treecopy(self.msa.h, self.t, src.t)
        self.root = src.root
        return self

    def arbreroot(self):
        path = np.zeros(self.msa.B, dtype=np.int64)
        k = C.c_long(0)
        r = self.L.lvbh_arbreroot(self.msa.h, self.t, self.root, path.ctypes.data, C.byref(k))
        if r < 0:
            raise HostError(self.L.lvbh_last_error().decode(errors="replace"))
        self.root = int(r)
        return self.root, path[: k.value].copy()

    def reroot(self, newroot: int) -> None:
        _check(self.L.lvbh_reroot(self.msa.h, self.t, self.root, newroot))
        self.root = newroot

    def get(self):
        B = self.msa.B
        parent, left, right = (np.zeros(B, dtype=np.int64) for _ in range(3))
        dirty = np.zeros(B, dtype=np.int32)
        self.L.lvbh_tree_get(self.msa.h, self.t, parent.ctypes.data, left.ctypes.data, right.ctypes.data,
                             dirty.ctypes.data)
        return parent, left, right, dirty

    def set_topology(self, parent, left, right, root: int) -> None:
        parent, left, right = (np.ascontiguousarray(a, dtype=np.int64) for a in (parent, left, right))
        self.L.lvbh_tree_set(self.msa.h, self.t, parent.ctypes.data, left.ctypes.data, right.ctypes.data)
        self.root = root

    def same_topology(self, other: "HostTree") -> bool:
        return bool(self.L.lvbh_same_topology(self.msa.h, self.t, self.root, other.t, other.root))

    def newick(self, path: str) -> None:
        _check(self.L.lvbh_treeprint(self.msa.h, self.t, self.root, path.encode()))

    def hillclimb(self, device: int = 0, outtree_path: Optional[str] = None) -> dict:
        """deterministic_hillclimb (Solve.c:64-142) from this tree, on the device"""
        out = (C.c_long * 4)()
        _check(self.L.lvbh_hillclimb(self.msa.h, device, self.t, self.root, outtree_path.encode() if outtree_path else None, out))
        return dict(length=int(out[0]), iterations=int(out[1]), trees=int(out[2]), neighbourhood_calls=int(out[3]))

    def getplen(self) -> int:
        """the compatible one-tree getplen (needs lvbh_device_open first)"""
        r = self.L.lvbh_getplen(self.msa.h, self.t, self.root)
        if r < 0:
            raise HostError(self.L.lvbh_last_error().decode(errors="replace"))
        return int(r)
