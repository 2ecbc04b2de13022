import sys
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parent.parent
if str(ROOT) not in sys.path:
    sys.path.insert(0, str(ROOT))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def oracle():
    """The CPU oracle (test infrastructure): built on first use."""
    import oracle as _oracle
    _oracle.build()
    return _oracle


@pytest.fixture(scope="module")
def cuda_env():
    """(device, stream) -- the counterpart of the reference's cl_env fixture
    (tests/conftest.py:4-12 upstream: (ctx, cq))."""
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    return dev, torch.cuda.Stream(dev)


@pytest.fixture(autouse=True)
def scratch_canaries():
    """COL_B200_DEBUG=1 (the -DCOL_DEBUG library: index / stack-depth traps in the kernels): after every
    test the canaries around each scratch buffer the package allocated are checked."""
    yield
    import os
    if os.environ.get("COL_B200_DEBUG") == "1":
        from collision_b200 import _lib
        _lib.check_canaries()
