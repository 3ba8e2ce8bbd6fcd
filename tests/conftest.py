import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (the parity tests proper; run on a B200)")


@pytest.fixture(scope="session")
def oracle():
    """The CPU parity oracle (oracle/libnimble_ref.so), built on demand.  Checker only."""
    from oracle.oracle_capi import oracle_backend
    return oracle_backend()


@pytest.fixture(scope="session")
def b200():
    """The product: nimble_b200/libnimb200.so through the C ABI.  No fallback: missing library = error."""
    import torch
    assert torch.cuda.is_available(), "gpu tests need a CUDA device"
    from nimble_b200 import _capi
    return _capi.default_backend()
