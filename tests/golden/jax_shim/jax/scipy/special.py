"""jax.scipy.special stand-in (see ../_core.py).  Tests only."""
import torch as _torch

from .._core import _t


def gammaln(x):
    return _torch.lgamma(_t(x))


def logsumexp(a, axis=None, b=None):
    assert b is None
    a = _t(a)
    return _torch.logsumexp(a, dim=tuple(range(a.ndim)) if axis is None else axis)
