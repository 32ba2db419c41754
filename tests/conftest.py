"""Shared pytest configuration.

Markers: ``gpu`` = needs a real B200 (the driver runs ``-m gpu`` on the GPU box and ``-m "not gpu"``
on the CPU-only box). GPU tests make sure the native library is built and loaded; there is no
PyTorch fallback for them to silently pass on.
"""
import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: test needs a CUDA device (B200)")
    config.addinivalue_line("markers", "multigpu: test needs at least 2 CUDA devices")


@pytest.fixture(scope="session")
def native_lib():
    from tiresias_b200 import build
    from tiresias_b200.ops import native
    build.build()
    return native.lib()


@pytest.fixture(scope="session")
def cuda_device(native_lib):
    import torch
    if not torch.cuda.is_available():
        pytest.fail("GPU test selected but no CUDA device is visible")
    torch.cuda.set_device(0)
    return torch.device("cuda", 0)
