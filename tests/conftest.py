import os
import sys
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parent.parent
PKG = ROOT / "scalable-hw-agnostic-inference_b200"
for p in (str(ROOT), str(PKG)):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (B200, sm_100a)")
    config.addinivalue_line("markers", "slow: minutes of host CPU time in the fp32 oracle (full-size, full-schedule parity)")


@pytest.fixture(scope="session")
def lib_built():
    """builds libb200sd.so if it is stale (nvcc cross-compiles without a GPU)"""
    sys.path.insert(0, str(PKG))
    import build as b200_build
    return b200_build.build()


@pytest.fixture(scope="session")
def cuda():
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    torch.backends.cuda.matmul.allow_tf32 = False
    torch.backends.cudnn.allow_tf32 = False
    return torch.device("cuda:0")
