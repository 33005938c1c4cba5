import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (B200, sm_100a); run with -m gpu on the GPU box")
    # fp32 references computed by torch on the GPU must be fp32: with TF32 on (cuDNN's default) F.conv2d drops 13
    # mantissa bits before it judges a kernel at 2e-3
    import torch
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False


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
