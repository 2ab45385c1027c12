"""Stand-in `jax` package for running the reference's source without jax (see _core.py).  Tests only."""
from ._core import jit, vmap, value_and_grad, grad, config, lax, random  # noqa: F401
from . import numpy  # noqa: F401
from . import scipy  # noqa: F401
import sys as _sys
_sys.modules[__name__ + ".random"] = random
_sys.modules[__name__ + ".lax"] = lax
Array = __import__("torch").Tensor
