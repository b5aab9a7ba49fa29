import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.dirname(os.path.abspath(__file__))):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")
    config.addinivalue_line("markers", "reference: needs /root/reference (build container only)")


@pytest.fixture(scope="session", autouse=True)
def _shim():
    # make `import pyccl` / `import healpy` resolve (real modules if present, stand-ins otherwise), so the
    # oracle and the product consume identical per-halo scalars
    from oracle import ref_shim
    ref_shim.install()
    yield


@pytest.fixture(scope="session")
def gpu():
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from baryonification_b200 import _lib
    _lib.load(require_cuda=True)  # must not silently fall back
    return torch.device("cuda", 0)
