"""pytest configuration: the `gpu` marker (tests that need a B200) and the repo root on sys.path."""
import os
import sys

import pytest

# Two contexts of one group on the SAME GPU (tests/test_gpu_multi.py on a one-GPU box) wait for each other on the device.
# With CUDA's default lazy module loading the first use of a kernel synchronises the device, which would wait for the other
# context's spinning exchange kernel: load every kernel at start-up instead.  Must be set before cuInit (below).
os.environ.setdefault("CUDA_MODULE_LOADING", "EAGER")

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (B200); run with -m gpu on the GPU box")


def _cuda_device_count():
    try:
        import ctypes
        cuda = ctypes.CDLL("libcuda.so.1")
        if cuda.cuInit(0) != 0:
            return 0
        n = ctypes.c_int(0)
        cuda.cuDeviceGetCount(ctypes.byref(n))
        return n.value
    except OSError:
        return 0


def pytest_collection_modifyitems(config, items):
    if _cuda_device_count() > 0:
        return
    skip = pytest.mark.skip(reason="no CUDA device in this container; run under gpurun with -m gpu")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)


@pytest.fixture(scope="session")
def oracle_lib():
    """The CPU oracle, built on demand (the checker — never the product path)."""
    from oracle import oracle
    oracle.build()
    return oracle
