import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


def _cuda_device_count():
    try:
        import ctypes
        cudart = ctypes.CDLL("libcudart.so")
    except OSError:
        try:
            import torch
            return torch.cuda.device_count()
        except Exception:
            return 0
    n = ctypes.c_int(0)
    rc = cudart.cudaGetDeviceCount(ctypes.byref(n))
    return n.value if rc == 0 else 0


def pytest_collection_modifyitems(config, items):
    have_gpu = None
    for item in items:
        if "gpu" in item.keywords:
            if have_gpu is None:
                have_gpu = _cuda_device_count() > 0
            if not have_gpu:
                item.add_marker(pytest.mark.skip(reason="no CUDA device"))
