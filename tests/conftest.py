import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: test needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def lib():
    import gol_b200
    return gol_b200._native.lib()


@pytest.fixture(scope="session")
def ref_serial():
    """The unmodified serial reference (src/game.c), built once; skip if unreachable."""
    from gol_b200.utils import reference
    exe = reference.build_serial()
    if exe is None:
        pytest.skip("reference sources not available")
    return exe


@pytest.fixture(scope="session")
def cuda_device():
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    return torch.device("cuda", 0)
