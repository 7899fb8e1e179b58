import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def cuda_lib():
    """Loads the C-ABI library and fails loudly (no skip, no fallback) when it cannot run."""
    import torch

    from omniisaacgymenvs_b200 import _lib

    if not torch.cuda.is_available():
        pytest.fail("a test marked gpu was run without a CUDA device")
    return _lib.load()
