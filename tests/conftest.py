import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


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
