import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


def pytest_sessionstart(session):
    """A fresh checkout has no built libraries (they are git-ignored): build them once, as __graft_entry__.build() does,
    so that the suite does not depend on the order in which the driver runs its checks."""
    from mcmc_b200 import _abi
    if all(os.path.exists(p) for p in _abi.LIB_PATHS.values()):
        return
    from mcmc_b200.csrc import build as cbuild
    if os.path.exists(cbuild.NVCC):
        cbuild.build(("fast", "strict"))


@pytest.fixture(scope="session")
def oracle():
    """The C oracle (test infrastructure), built on demand."""
    from oracle import oracle as O
    O.build()
    O.lib()
    return O


@pytest.fixture(scope="session")
def gpu_lib():
    from mcmc_b200 import _abi
    L = _abi.lib("fast")
    if L.mcmc_device_count() < 1:
        pytest.fail("-m gpu test without a CUDA device: the product path has no CPU fallback")
    return L
