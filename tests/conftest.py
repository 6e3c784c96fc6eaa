import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def cuda():
    """Loads the C-ABI library on a CUDA device; GPU tests fail loudly if it is missing."""
    import torch
    assert torch.cuda.is_available(), "GPU test started without a CUDA device"
    from theia_b200 import _lib
    _lib.load()
    return torch.device("cuda")


@pytest.fixture(scope="session")
def orc():
    import oracle
    oracle.build()
    return oracle
