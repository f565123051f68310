"""ctypes binding of include/lvb_gpu.h (lvb_b200/lib/liblvbgpu.so).

Plumbing only: every call goes straight to the C-ABI; there is no arithmetic and no CPU
fallback here.  If the library is missing, import fails loudly.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Iterable, Optional, Sequence

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "lib", "liblvbgpu.so")


class GpuError(RuntimeError):
    """A non-zero status from the C-ABI (LVB host code would crash() on it, src/Error.c:49)."""


class _Topo(C.Structure):
    _fields_ = [
        ("left", C.c_void_p),
        ("right", C.c_void_p),
        ("stride_bytes", C.c_int64),
        ("root", C.c_int64),
        ("dirty", C.c_void_p),
        ("n_dirty", C.c_int64),
        ("mod_node", C.c_void_p),
        ("mod_left", C.c_void_p),
        ("mod_right", C.c_void_p),
        ("n_mod", C.c_int64),
    ]


_lib = None


def load_library() -> C.CDLL:
    """Load liblvbgpu.so; raises if it has not been built (no fallback)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} not found: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "or `make -C lvb_b200/csrc`. lvb_b200 has no CPU fallback."
        )
    lib = C.CDLL(LIB_PATH)
    vp, i64, i32p = C.c_void_p, C.c_int64, C.POINTER(C.c_int)
    lib.lvb_gpu_last_error.restype = C.c_char_p
    lib.lvb_gpu_device_count.argtypes = [i32p]
    lib.lvb_gpu_init.argtypes = [C.c_int, C.POINTER(vp)]
    lib.lvb_gpu_shutdown.argtypes = [vp]
    lib.lvb_gpu_stream.restype = vp
    lib.lvb_gpu_stream.argtypes = [vp]
    lib.lvb_gpu_matrix_upload.argtypes = [vp, i64, i64, C.POINTER(vp), i64, i64, i64]
    lib.lvb_gpu_matrix_upload_flat.argtypes = [vp, i64, i64, vp, i64, i64, i64]
    lib.lvb_gpu_matrix_pack.argtypes = [vp, i64, i64, C.POINTER(C.c_char_p), i64, i64, i64]
    lib.lvb_gpu_column_stats.argtypes = [vp, i64, i64, C.POINTER(C.c_char_p), vp, C.POINTER(i64), C.POINTER(i64), C.POINTER(i64)]
    lib.lvb_gpu_tree_create.argtypes = [vp, i32p]
    lib.lvb_gpu_tree_destroy.argtypes = [vp, C.c_int]
    lib.lvb_gpu_tree_copy.argtypes = [vp, C.c_int, C.c_int]
    lib.lvb_gpu_tree_commit.argtypes = [vp, C.c_int, C.POINTER(_Topo), C.POINTER(i64)]
    lib.lvb_gpu_commit_batch.argtypes = [vp, i64, i32p, C.POINTER(_Topo), C.POINTER(i64)]
    lib.lvb_gpu_score_batch.argtypes = [vp, i64, i32p, C.POINTER(_Topo), C.POINTER(i64)]
    lib.lvb_gpu_score_batch_submit.argtypes = [vp, i64, i32p, C.POINTER(_Topo), i32p]
    lib.lvb_gpu_batch_wait.argtypes = [vp, C.c_int, C.POINTER(i64)]
    lib.lvb_gpu_score_keep.argtypes = [vp, i64, i32p, C.POINTER(_Topo), C.POINTER(C.c_uint8), C.POINTER(i64), i32p]
    lib.lvb_gpu_accept_kept.argtypes = [vp, C.c_int, i64, C.POINTER(i64)]
    lib.lvb_gpu_commit_batch_submit.argtypes = [vp, i64, i32p, C.POINTER(_Topo), i32p]
    lib.lvb_gpu_nni_neighbourhood.argtypes = [vp, C.c_int, C.c_int, C.POINTER(_Topo), C.POINTER(i64)]
    lib.lvb_gpu_download_row.argtypes = [vp, C.c_int, i64, vp]
    lib.lvb_gpu_tree_changes.argtypes = [vp, C.c_int, vp]
    lib.lvb_gpu_comm_unique_id.argtypes = [vp]
    lib.lvb_gpu_comm_init.argtypes = [vp, C.c_int, C.c_int, vp]
    lib.lvb_gpu_set_timing.argtypes = [vp, C.c_int]
    lib.lvb_gpu_launch_count.restype = i64
    lib.lvb_gpu_launch_count.argtypes = [vp]
    lib.lvb_gpu_last_kernel_ms.argtypes = [vp, C.POINTER(C.c_float)]
    lib.lvb_gpu_kernel_time.argtypes = [vp, C.POINTER(C.c_double), C.POINTER(i64), C.c_int]
    lib.lvb_gpu_rerun_last_batch.argtypes = [vp, C.POINTER(i64)]
    lib.lvb_gpu_synchronize.argtypes = [vp]
    lib.lvb_gpu_compile_schedule.argtypes = [i64, C.POINTER(_Topo), C.c_int, vp, i64, C.POINTER(i64), C.POINTER(i64)]
    lib.lvb_gpu_compile_pieces.argtypes = [i64, C.POINTER(_Topo), C.c_int, vp, i64, C.POINTER(i64), C.POINTER(i64), vp, i64, C.POINTER(i64)]
    lib.lvb_gpu_compile_up_pass.argtypes = [i64, C.POINTER(_Topo), C.c_int, vp, i64, C.POINTER(i64), C.POINTER(i64), vp, i64, C.POINTER(i64)]
    _lib = lib
    return lib


def _check(rc: int) -> None:
    if rc != 0:
        raise GpuError(load_library().lvb_gpu_last_error().decode(errors="replace"))


class Topology:
    """A (candidate) tree as getplen sees it: child links, root, dirty internal nodes.

    ``left``/``right`` are int64 arrays over all 2n-3 nodes (leaves hold -1, LVB's UNSET);
    ``dirty`` is None for "every internal node".  Arrays are kept alive by this object.
    """

    def __init__(self, left, right, root: int, dirty: Optional[Iterable[int]] = None, overlay=None):
        self.left = np.ascontiguousarray(left, dtype=np.int64)
        self.right = np.ascontiguousarray(right, dtype=np.int64)
        self.root = int(root)
        # overlay = (nodes, lefts, rights): child links that replace those of left/right
        self.overlay = None if overlay is None else tuple(np.ascontiguousarray(a, dtype=np.int64) for a in overlay)
        self.dirty = None if dirty is None else np.ascontiguousarray(list(dirty) if not isinstance(dirty, np.ndarray) else dirty, dtype=np.int64)

    def fill(self, t: _Topo) -> None:
        t.left = self.left.ctypes.data
        t.right = self.right.ctypes.data
        t.stride_bytes = 8
        t.root = self.root
        if self.dirty is None:
            t.dirty = None
            t.n_dirty = -1
        else:
            t.dirty = self.dirty.ctypes.data if self.dirty.size else None
            t.n_dirty = int(self.dirty.size)
        if self.overlay is None or self.overlay[0].size == 0:
            t.mod_node = t.mod_left = t.mod_right = None
            t.n_mod = 0
        else:
            t.mod_node, t.mod_left, t.mod_right = (a.ctypes.data for a in self.overlay)
            t.n_mod = int(self.overlay[0].size)

    def as_struct(self) -> _Topo:
        t = _Topo()
        self.fill(t)
        return t


def compile_schedule(n: int, topo: Topology, store: bool = False):
    """Host-only: the instruction stream for `topo` as an (k, 4) int32 array and its slot count."""
    lib = load_library()
    ts = topo.as_struct()
    k, slots = C.c_int64(0), C.c_int64(0)
    _check(lib.lvb_gpu_compile_schedule(n, C.byref(ts), int(store), None, 0, C.byref(k), C.byref(slots)))
    out = np.zeros((k.value, 4), dtype=np.int32)
    _check(lib.lvb_gpu_compile_schedule(n, C.byref(ts), int(store), out.ctypes.data, k.value, C.byref(k), C.byref(slots)))
    return out, int(slots.value)


def compile_up_pass(n: int, topo: Topology, piece_size: int = 0):
    """Host-only: the up-pass program of the NNI neighbourhood: (instr (k,4) int32, slots, pieces (p,3)
    int32 rows {instr_begin, n_instr, level})."""
    lib = load_library()
    ts = topo.as_struct()
    k, slots, npc = C.c_int64(0), C.c_int64(0), C.c_int64(0)
    _check(lib.lvb_gpu_compile_up_pass(n, C.byref(ts), piece_size, None, 0, C.byref(k), C.byref(slots), None, 0, C.byref(npc)))
    out = np.zeros((k.value, 4), dtype=np.int32)
    pieces = np.zeros((npc.value, 3), dtype=np.int32)
    _check(lib.lvb_gpu_compile_up_pass(n, C.byref(ts), piece_size, out.ctypes.data, k.value, C.byref(k), C.byref(slots),
                                       pieces.ctypes.data, npc.value, C.byref(npc)))
    return out, int(slots.value), pieces


def compile_pieces(n: int, topo: Topology, piece_size: int):
    """Host-only: a store program cut into pieces: (instr (k,4) int32, pieces (p,3) int32
    rows {instr_begin, n_instr, level}, slots)."""
    lib = load_library()
    ts = topo.as_struct()
    k, slots, npc = C.c_int64(0), C.c_int64(0), C.c_int64(0)
    _check(lib.lvb_gpu_compile_pieces(n, C.byref(ts), piece_size, None, 0, C.byref(k), C.byref(slots), None, 0, C.byref(npc)))
    instr = np.zeros((k.value, 4), dtype=np.int32)
    pieces = np.zeros((npc.value, 3), dtype=np.int32)
    _check(lib.lvb_gpu_compile_pieces(n, C.byref(ts), piece_size, instr.ctypes.data, k.value, C.byref(k), C.byref(slots),
                                      pieces.ctypes.data, npc.value, C.byref(npc)))
    return instr, pieces, int(slots.value)


class Engine:
    """One context = one GPU, one stream (lvb_gpu_init .. lvb_gpu_shutdown)."""

    def __init__(self, device: int = 0):
        self.lib = load_library()
        self.ctx = C.c_void_p()
        _check(self.lib.lvb_gpu_init(device, C.byref(self.ctx)))
        _check(self.lib.lvb_gpu_set_timing(self.ctx, 1))  # tests and bench read kernel times
        self.n = self.nwords = self.word_count = 0

    def close(self) -> None:
        if self.ctx:
            self.lib.lvb_gpu_shutdown(self.ctx)
            self.ctx = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- alignment ----
    def upload(self, enc: np.ndarray, word_begin: int = 0, word_count: Optional[int] = None, max_trees: int = 1) -> None:
        enc = np.ascontiguousarray(enc, dtype=np.uint64)
        n, nwords = enc.shape
        wc = nwords - word_begin if word_count is None else word_count
        _check(self.lib.lvb_gpu_matrix_upload_flat(self.ctx, n, nwords, enc.ctypes.data, word_begin, wc, max_trees))
        self.n, self.nwords, self.word_count = n, nwords, wc

    def upload_rows(self, rows: Sequence[np.ndarray], word_begin: int = 0, word_count: Optional[int] = None, max_trees: int = 1) -> None:
        """Non-contiguous rows, the layout GetSoln builds (Solve.c:516-519)."""
        rows = [np.ascontiguousarray(r, dtype=np.uint64) for r in rows]
        n, nwords = len(rows), rows[0].size
        ptrs = (C.c_void_p * n)(*[r.ctypes.data for r in rows])
        wc = nwords - word_begin if word_count is None else word_count
        _check(self.lib.lvb_gpu_matrix_upload(self.ctx, n, nwords, ptrs, word_begin, wc, max_trees))
        self.n, self.nwords, self.word_count = n, nwords, wc

    def pack(self, rows: Sequence[bytes], word_begin: int = 0, word_count: Optional[int] = None, max_trees: int = 1) -> None:
        n, m = len(rows), len(rows[0])
        nwords = (m + 15) // 16
        arr = (C.c_char_p * n)(*rows)
        wc = nwords - word_begin if word_count is None else word_count
        _check(self.lib.lvb_gpu_matrix_pack(self.ctx, n, m, arr, word_begin, wc, max_trees))
        self.n, self.nwords, self.word_count = n, nwords, wc

    def column_stats(self, rows: Sequence[bytes]):
        """(keep[m] uint8, variable columns, MinimumTreeLength of all columns, of the variable columns)"""
        n, m = len(rows), len(rows[0])
        arr = (C.c_char_p * n)(*rows)
        keep = np.zeros(m, dtype=np.uint8)
        a, b, c_ = C.c_int64(0), C.c_int64(0), C.c_int64(0)
        _check(self.lib.lvb_gpu_column_stats(self.ctx, n, m, arr, keep.ctypes.data, C.byref(a), C.byref(b), C.byref(c_)))
        return keep, int(a.value), int(b.value), int(c_.value)

    # ---- trees ----
    def tree_create(self) -> int:
        t = C.c_int(-1)
        _check(self.lib.lvb_gpu_tree_create(self.ctx, C.byref(t)))
        return t.value

    def tree_destroy(self, tree: int) -> None:
        _check(self.lib.lvb_gpu_tree_destroy(self.ctx, tree))

    def tree_copy(self, dst: int, src: int) -> None:
        _check(self.lib.lvb_gpu_tree_copy(self.ctx, dst, src))

    def commit(self, tree: int, topo: Topology, wait: bool = True) -> Optional[int]:
        ts = topo.as_struct()
        if not wait:
            _check(self.lib.lvb_gpu_tree_commit(self.ctx, tree, C.byref(ts), None))
            return None
        length = C.c_int64(0)
        _check(self.lib.lvb_gpu_tree_commit(self.ctx, tree, C.byref(ts), C.byref(length)))
        return int(length.value)

    def _batch(self, fn, trees: Sequence[int], topos: Sequence[Topology]) -> np.ndarray:
        k = len(topos)
        tarr = (C.c_int * k)(*[int(t) for t in trees])
        sarr = (_Topo * k)()
        for i, tp in enumerate(topos):
            tp.fill(sarr[i])
        out = np.zeros(k, dtype=np.int64)
        _check(fn(self.ctx, k, tarr, sarr, out.ctypes.data_as(C.POINTER(C.c_int64))))
        return out

    def commit_batch(self, trees: Sequence[int], topos: Sequence[Topology]) -> np.ndarray:
        return self._batch(self.lib.lvb_gpu_commit_batch, trees, topos)

    def score_batch(self, trees: Sequence[int], topos: Sequence[Topology]) -> np.ndarray:
        return self._batch(self.lib.lvb_gpu_score_batch, trees, topos)

    def prepare_batch(self, trees: Sequence[int], topos: Sequence[Topology]):
        """the C arrays of a batch (device tree numbers, lvb_gpu_topology records pointing into the arrays of
        `topos`, which must stay alive): what a C caller holds anyway; score_prepared then times the C-ABI call
        itself, not this binding"""
        k = len(topos)
        tarr = (C.c_int * k)(*[int(t) for t in trees])
        sarr = (_Topo * k)()
        for i, tp in enumerate(topos):
            tp.fill(sarr[i])
        return k, tarr, sarr, np.zeros(k, dtype=np.int64), list(topos)

    def score_prepared(self, batch) -> np.ndarray:
        k, tarr, sarr, out, _ = batch
        _check(self.lib.lvb_gpu_score_batch(self.ctx, k, tarr, sarr, out.ctypes.data_as(C.POINTER(C.c_int64))))
        return out

    def score_batch_submit(self, trees: Sequence[int], topos: Sequence[Topology]):
        """enqueue a score batch; returns (ticket, n) for batch_wait"""
        k = len(topos)
        tarr = (C.c_int * k)(*[int(t) for t in trees])
        sarr = (_Topo * k)()
        for i, tp in enumerate(topos):
            tp.fill(sarr[i])
        ticket = C.c_int(-1)
        _check(self.lib.lvb_gpu_score_batch_submit(self.ctx, k, tarr, sarr, C.byref(ticket)))
        return ticket.value, k

    def commit_batch_submit(self, trees: Sequence[int], topos: Sequence[Topology]):
        """enqueue a commit batch; returns (ticket, n) for batch_wait (the topologies must stay alive
        only for the duration of this call: they are compiled here)"""
        k = len(topos)
        tarr = (C.c_int * k)(*[int(t) for t in trees])
        sarr = (_Topo * k)()
        for i, tp in enumerate(topos):
            tp.fill(sarr[i])
        ticket = C.c_int(-1)
        _check(self.lib.lvb_gpu_commit_batch_submit(self.ctx, k, tarr, sarr, C.byref(ticket)))
        return ticket.value, k

    def score_keep(self, trees: Sequence[int], topos: Sequence[Topology], keep: Sequence[int]):
        """lvb_gpu_score_keep: lengths of the candidates; those with keep[k] != 0 leave their rows in the spare
        bank of their tree.  Returns (lengths, ticket) — ticket for accept_kept"""
        k = len(topos)
        tarr = (C.c_int * k)(*[int(t) for t in trees])
        sarr = (_Topo * k)()
        for i, tp in enumerate(topos):
            tp.fill(sarr[i])
        karr = (C.c_uint8 * k)(*[1 if x else 0 for x in keep])
        out = np.zeros(k, dtype=np.int64)
        ticket = C.c_int(-1)
        _check(self.lib.lvb_gpu_score_keep(self.ctx, k, tarr, sarr, karr, out.ctypes.data_as(C.POINTER(C.c_int64)),
                                           C.byref(ticket)))
        return out, ticket.value

    def accept_kept(self, ticket: int, cands: Sequence[int]) -> None:
        """lvb_gpu_accept_kept: the kept candidates `cands` of the batch become their trees"""
        arr = (C.c_int64 * len(cands))(*[int(x) for x in cands])
        _check(self.lib.lvb_gpu_accept_kept(self.ctx, int(ticket), len(cands), arr))

    def batch_wait(self, ticket) -> np.ndarray:
        out = np.zeros(ticket[1], dtype=np.int64)
        _check(self.lib.lvb_gpu_batch_wait(self.ctx, ticket[0], out.ctypes.data_as(C.POINTER(C.c_int64))))
        return out

    def nni_neighbourhood(self, tree: int, scratch_tree: int, topo: Topology) -> np.ndarray:
        """lengths of all 2(n-3) mutate_deterministic candidates of the evaluated `tree`
        (index 2*(u-n)+left); `scratch_tree` receives the up sets"""
        ts = topo.as_struct()
        out = np.zeros(2 * (self.n - 3), dtype=np.int64)
        _check(self.lib.lvb_gpu_nni_neighbourhood(self.ctx, tree, scratch_tree, C.byref(ts),
                                                  out.ctypes.data_as(C.POINTER(C.c_int64))))
        return out

    # ---- parity helpers ----
    def download_row(self, tree: int, node: int) -> np.ndarray:
        out = np.zeros(self.word_count, dtype=np.uint64)
        _check(self.lib.lvb_gpu_download_row(self.ctx, tree, node, out.ctypes.data))
        return out

    def changes(self, tree: int) -> np.ndarray:
        out = np.zeros(2 * self.n - 3, dtype=np.int64)
        _check(self.lib.lvb_gpu_tree_changes(self.ctx, tree, out.ctypes.data))
        return out

    # ---- multi-GPU ----
    @staticmethod
    def comm_unique_id() -> bytes:
        buf = C.create_string_buffer(128)
        _check(load_library().lvb_gpu_comm_unique_id(buf))
        return buf.raw

    def comm_init(self, nranks: int, rank: int, uid: bytes) -> None:
        buf = C.create_string_buffer(uid, 128)
        _check(self.lib.lvb_gpu_comm_init(self.ctx, nranks, rank, buf))

    # ---- introspection ----
    @property
    def launches(self) -> int:
        return int(self.lib.lvb_gpu_launch_count(self.ctx))

    def last_kernel_ms(self) -> float:
        ms = C.c_float(0)
        _check(self.lib.lvb_gpu_last_kernel_ms(self.ctx, C.byref(ms)))
        return float(ms.value)

    def kernel_time(self, reset: bool = False):
        """(total ms, launches) of the Fitch kernel since the last reset (CUDA events on the stream)"""
        ms, k = C.c_double(0), C.c_int64(0)
        _check(self.lib.lvb_gpu_kernel_time(self.ctx, C.byref(ms), C.byref(k), int(reset)))
        return float(ms.value), int(k.value)

    def rerun_last_batch(self, count: int = 0) -> Optional[np.ndarray]:
        """launch the device-resident batch again; count > 0 -> wait and return its lengths"""
        if count <= 0:
            _check(self.lib.lvb_gpu_rerun_last_batch(self.ctx, None))
            return None
        out = np.zeros(count, dtype=np.int64)
        _check(self.lib.lvb_gpu_rerun_last_batch(self.ctx, out.ctypes.data_as(C.POINTER(C.c_int64))))
        return out

    def synchronize(self) -> None:
        _check(self.lib.lvb_gpu_synchronize(self.ctx))

    @property
    def stream(self) -> int:
        return int(self.lib.lvb_gpu_stream(self.ctx) or 0)
