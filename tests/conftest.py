import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a real B200 (run with -m gpu on the GPU box)")


@pytest.fixture(scope="session")
def built():
    """Build (or reuse) the product library and the checkers."""
    import __graft_entry__ as g
    g.build()
    return True


@pytest.fixture(scope="session")
def gpu(built):
    import torch
    from darknet_b200 import lib as L
    assert torch.cuda.is_available(), "GPU test selected but no CUDA device (no CPU fallback exists)"
    L.check(L.lib().yc_set_device(0), "yc_set_device")
    torch.cuda.set_device(0)
    return torch.device("cuda:0")
