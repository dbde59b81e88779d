import os
import sys

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
sys.dont_write_bytecode = True


# the torch-side ConvEncode / ConvDecode must be fp32 for parity against the CPU oracle (cuDNN defaults to TF32)
torch.backends.cudnn.allow_tf32 = False
torch.backends.cuda.matmul.allow_tf32 = False


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


def pytest_collection_modifyitems(config, items):
    if torch.cuda.is_available():
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for it in items:
        if "gpu" in it.keywords:
            it.add_marker(skip)


@pytest.fixture
def fake_backend():
    """Swap the CUDA backend for the torch-CPU emulation in tests/fake_backend.py (host-logic tests only)."""
    from ode_rl_b200 import kernels
    from tests.fake_backend import FakeBackend

    fb = FakeBackend()
    prev = kernels.set_backend(fb)
    try:
        yield fb
    finally:
        kernels.set_backend(prev)


@pytest.fixture(scope="session")
def reference():
    """The real reference modules (only where /root/reference is mounted: the build container)."""
    from oracle import refload

    if not refload.available():
        pytest.skip("reference tree not mounted")
    return refload.load()
