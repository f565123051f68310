"""GPU: lvb_gpu_column_stats_noalloc — the allocation-free variant of lvb_gpu_column_stats (scratch in a fixed
device array inside the library, no counterpart in the reference) —
gives exactly the numbers of the existing entry point, on calls in a row that share the scratch array.

Collected last on purpose (file name).  The CALLER's device buffers come from the
CUDA runtime library through ctypes (test-side plumbing only; the entry point itself allocates nothing) — not from
torch: importing torch late in a process in which the engine has already loaded the system's NCCL fails on a
symbol of torch's newer NCCL, and this test must also pass after all the others."""
import ctypes as C
import glob
import os

import numpy as np
import pytest

from lvb_b200 import Engine, load_library

pytestmark = pytest.mark.gpu

MAX_COLUMNS = 4194304  # LVB_GPU_NOALLOC_MAX_COLUMNS, include/lvb_gpu.h


def _alignment(n, m, seed):
    rng = np.random.default_rng(seed)
    sym = np.frombuffer(b"ACGTUYRWSKMBDHVNX?O-", dtype=np.uint8)
    t = sym[rng.integers(0, len(sym), size=(n, m))].copy()
    t[:, rng.random(m) < 0.2] = ord("A")  # constant columns
    t[:, 5] = ord("-")                     # no countable character: adds -1
    t[:, 7] = ord("N")
    t[n // 2, 7] = ord("X")                # variable, but nothing countable
    return t


def _bind():
    lib = load_library()
    lib.lvb_gpu_column_stats_noalloc.argtypes = [C.c_void_p, C.c_int64, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p]
    lib.lvb_gpu_column_stats_noalloc.restype = C.c_int
    return lib


_rt = None


def _cudart():
    """the CUDA runtime as a shared library (same primary context as the engine's statically linked one)"""
    global _rt
    if _rt is not None:
        return _rt
    cands = ["libcudart.so.12", "libcudart.so"] + sorted(glob.glob("/usr/local/cuda*/lib64/libcudart.so*"))
    try:
        import nvidia.cuda_runtime  # the wheel torch depends on; a plain package, importing it loads nothing

        cands += sorted(glob.glob(os.path.join(os.path.dirname(nvidia.cuda_runtime.__file__), "lib", "libcudart.so*")))
    except Exception:
        pass
    for name in cands:
        try:
            rt = C.CDLL(name)
        except OSError:
            continue
        rt.cudaMalloc.argtypes = [C.POINTER(C.c_void_p), C.c_size_t]
        rt.cudaMemcpy.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_int]
        rt.cudaMemset.argtypes = [C.c_void_p, C.c_int, C.c_size_t]
        rt.cudaFree.argtypes = [C.c_void_p]
        assert rt.cudaSetDevice(0) == 0
        _rt = rt
        return rt
    pytest.fail("no libcudart found for the caller-side buffers of this test")


class _DeviceBuffer:
    def __init__(self, nbytes, fill=None):
        self.rt, self.nbytes, self.ptr = _cudart(), nbytes, C.c_void_p()
        assert self.rt.cudaMalloc(C.byref(self.ptr), max(1, nbytes)) == 0
        if fill is not None:
            assert self.rt.cudaMemset(self.ptr, fill, nbytes) == 0

    def upload(self, arr):
        arr = np.ascontiguousarray(arr)
        assert arr.nbytes == self.nbytes
        assert self.rt.cudaMemcpy(self.ptr, arr.ctypes.data, arr.nbytes, 1) == 0  # cudaMemcpyHostToDevice
        return self

    def download(self, dtype):
        out = np.zeros(self.nbytes // np.dtype(dtype).itemsize, dtype=dtype)
        assert self.rt.cudaMemcpy(out.ctypes.data, self.ptr, self.nbytes, 2) == 0  # cudaMemcpyDeviceToHost
        return out

    def free(self):
        self.rt.cudaFree(self.ptr)


def _noalloc(lib, eng, text, with_keep=True):
    n, m = text.shape
    d_text = _DeviceBuffer(n * m).upload(text)
    d_keep = _DeviceBuffer(m, fill=7)
    d_out = _DeviceBuffer(24, fill=0x5A)
    try:
        assert _cudart().cudaDeviceSynchronize() == 0  # the entry point works on the context's own stream
        rc = lib.lvb_gpu_column_stats_noalloc(eng.ctx, n, m, d_text.ptr, d_keep.ptr if with_keep else None, d_out.ptr)
        assert rc == 0, lib.lvb_gpu_last_error()
        eng.synchronize()
        return d_keep.download(np.uint8), [int(x) for x in d_out.download(np.int64)]
    finally:
        for b in (d_text, d_keep, d_out):
            b.free()


def test_noalloc_matches_column_stats_on_calls_in_a_row():
    lib = _bind()
    eng = Engine(0)
    try:
        launches = eng.launches
        # shapes: several CTAs, one CTA, a ragged last CTA; the second and third call find the first one's partial
        # sums in the scratch array
        for call, (n, m, seed) in enumerate([(37, 5003, 1), (11, 200, 2), (64, 70_001, 3), (37, 5003, 1)]):
            text = _alignment(n, m, seed)
            keep, kept, all_len, kept_len = eng.column_stats([bytes(r) for r in text])
            got_keep, got = _noalloc(lib, eng, text)
            assert got == [kept, all_len, kept_len], (call, got, (kept, all_len, kept_len))
            assert np.array_equal(got_keep, keep), call
            assert 0 < kept < m
            # keep is optional
            _, got2 = _noalloc(lib, eng, text, with_keep=False)
            assert got2 == got, call
        assert eng.launches > launches
    finally:
        eng.close()


def test_noalloc_refuses_what_the_scratch_cannot_hold():
    lib = _bind()
    eng = Engine(0)
    buf = _DeviceBuffer(512, fill=0)
    try:
        rc = lib.lvb_gpu_column_stats_noalloc(eng.ctx, 4, MAX_COLUMNS + 1, buf.ptr, None, buf.ptr)
        assert rc != 0 and b"LVB_GPU_NOALLOC_MAX_COLUMNS" in lib.lvb_gpu_last_error()
        rc = lib.lvb_gpu_column_stats_noalloc(eng.ctx, 4, 16, None, None, buf.ptr)
        assert rc != 0 and b"null argument" in lib.lvb_gpu_last_error()
    finally:
        buf.free()
        eng.close()
