import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (B200, sm_100a); run with -m gpu on the GPU box")


@pytest.fixture(scope="session")
def b200sd():
    import b200sd as pkg
    return pkg


@pytest.fixture(scope="session")
def cuda_lib(b200sd):
    """The loaded C-ABI library on a GPU box; fails loudly (no fallback) if missing."""
    import torch
    assert torch.cuda.is_available(), "GPU tests need CUDA"
    b200sd.lib.load()
    info = b200sd.lib.device_info()
    assert info["cc"][0] == 10, info
    return b200sd.lib
