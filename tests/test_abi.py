"""CPU: the C-ABI library loads without a GPU and exports every symbol include/lvb_gpu.h
declares; calls that need a device fail loudly (no CPU fallback)."""
import ctypes as C
import os
import re

import pytest

from lvb_b200 import gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    h = open(os.path.join(ROOT, "include", "lvb_gpu.h")).read()
    h = re.sub(r"/\*.*?\*/", "", h, flags=re.S)
    return sorted(set(re.findall(r"\b(lvb_gpu_[a-z0-9_]+)\s*\(", h)))


def test_header_symbols_are_exported():
    lib = gpu.load_library()
    names = declared_symbols()
    assert len(names) >= 18
    for name in names:
        assert hasattr(lib, name), f"{name} declared in include/lvb_gpu.h but not exported"


def test_no_torch_types_in_header():
    h = open(os.path.join(ROOT, "include", "lvb_gpu.h")).read()
    assert 'extern "C"' in h
    h = re.sub(r"/\*.*?\*/", "", h, flags=re.S)
    assert "torch" not in h.lower() and "at::" not in h and "std::" not in h


def test_no_cpu_fallback_without_device():
    lib = gpu.load_library()
    n = C.c_int(0)
    rc = lib.lvb_gpu_device_count(C.byref(n))
    if rc == 0 and n.value > 0:
        pytest.skip("a GPU is present")
    with pytest.raises(gpu.GpuError):
        gpu.Engine(0)


def test_product_does_not_touch_oracle():
    """nothing under lvb_b200/ may import, link or load oracle/"""
    for dirpath, _, files in os.walk(os.path.join(ROOT, "lvb_b200")):
        for f in files:
            if f.endswith((".py", ".cu", ".cpp", ".h", ".cuh", "Makefile")):
                text = open(os.path.join(dirpath, f), errors="replace").read()
                assert "liboracle" not in text and "libLVB_ref" not in text and "oracle/" not in text.replace(
                    "has no CPU fallback", ""), f
