import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a real B200 (run with -m gpu on the GPU box)")


def _have_gpu() -> bool:
    try:
        import ctypes

        from lvb_b200 import load_library

        n = ctypes.c_int(0)
        return load_library().lvb_gpu_device_count(ctypes.byref(n)) == 0 and n.value > 0
    except Exception:
        return False


def pytest_collection_modifyitems(config, items):
    # -m gpu tests must FAIL, not skip, when the extension cannot reach a GPU on the GPU box;
    # they are only deselected by the driver's -m "not gpu" run here.
    pass


@pytest.fixture(scope="session")
def golden_dir():
    return os.path.join(ROOT, "tests", "golden")


@pytest.fixture(scope="session")
def engine():
    from lvb_b200 import Engine

    e = Engine(0)
    yield e
    e.close()
