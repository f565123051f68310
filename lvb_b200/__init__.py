"""lvb_b200 — B200 (sm_100a) Fitch tree-length engine for LVB.

The product is the C-ABI shared library ``lvb_b200/lib/liblvbgpu.so`` declared in
``include/lvb_gpu.h`` (CUDA kernels + host schedule compiler) and the C++ host layer
``lvb_b200/lib/liblvbhost.so`` that mirrors LVB's own function names on top of it.  This
Python package is a thin ctypes binding over that C-ABI used by the tests and bench.py; it
contains no arithmetic and has no CPU fallback: importing :mod:`lvb_b200.gpu` raises if the
library has not been built.
"""

from .gpu import Engine, GpuError, Topology, load_library, compile_schedule, compile_pieces, compile_up_pass  # noqa: F401
