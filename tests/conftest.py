import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (B200); run with -m gpu")


@pytest.fixture(scope="session")
def lib_built():
    """Make sure libjtsm_b200.so exists (nvcc cross-compiles without a GPU)."""
    from jtsm_b200 import build

    return build.build_library()


@pytest.fixture(scope="session")
def oracle():
    from oracle import capi

    capi.lib()
    return capi


@pytest.fixture(scope="session")
def cuda():
    import torch

    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    return torch.device("cuda:0")
