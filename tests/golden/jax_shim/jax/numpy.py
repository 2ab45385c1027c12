"""jax.numpy stand-in on torch.float64 CPU tensors (see _core.py).  Tests only."""
import builtins as _b
import math as _math

import torch as _torch

from ._core import _t, _dt

ndarray = _torch.Tensor
float64, float32, int32, int64, bool_ = _torch.float64, _torch.float32, _torch.int32, _torch.int64, _torch.bool
pi, inf, nan = _math.pi, _math.inf, _math.nan


def asarray(x, dtype=None):
    return _t(x, dtype)


def array(x, dtype=None):
    if isinstance(x, (list, tuple)) and _b.any(isinstance(v, _torch.Tensor) for v in x):
        return _torch.stack([_t(v) for v in x]).to(_dt(dtype) or _torch.float64)
    return _t(x, dtype).clone()


def _shape(s):
    return (int(s),) if not isinstance(s, (tuple, list)) else tuple(int(v) for v in s)


def zeros(shape, dtype=float64): return _torch.zeros(_shape(shape), dtype=_dt(dtype))
def ones(shape, dtype=float64): return _torch.ones(_shape(shape), dtype=_dt(dtype))
def eye(n, dtype=float64): return _torch.eye(int(n), dtype=_dt(dtype))
def arange(*a, dtype=None): return _torch.arange(*a, dtype=_dt(dtype))
def linspace(a, b, n, dtype=float64): return _torch.linspace(float(a), float(b), int(n), dtype=_dt(dtype))
def diag(x): return _torch.diag(_t(x))
def concatenate(xs, axis=0): return _torch.cat([_t(x) for x in xs], dim=axis)
def stack(xs, axis=0): return _torch.stack([_t(x) for x in xs], dim=axis)
def atleast_1d(x): return _torch.atleast_1d(_t(x))
def atleast_2d(x): return _torch.atleast_2d(_t(x))
def log(x): return _torch.log(_t(x))
def exp(x): return _torch.exp(_t(x))
def sqrt(x): return _torch.sqrt(_t(x))
def tanh(x): return _torch.tanh(_t(x))
def sin(x): return _torch.sin(_t(x))
def cos(x): return _torch.cos(_t(x))
def abs(x): return _torch.abs(_t(x))
def square(x): return _t(x) ** 2
def isfinite(x): return _torch.isfinite(_t(x))
def isnan(x): return _torch.isnan(_t(x))
def isinf(x): return _torch.isinf(_t(x))
def dot(a, b): return _torch.dot(_t(a), _t(b)) if _t(a).ndim == 1 and _t(b).ndim == 1 else _t(a) @ _t(b)


def _red(fn, x, axis, keepdims=False):
    x = _t(x)
    if axis is None:
        return fn(x)
    r = fn(x, dim=axis, keepdim=keepdims)
    return r if isinstance(r, _torch.Tensor) else r[0]


def sum(x, axis=None, keepdims=False): return _red(_torch.sum, x, axis, keepdims)
def mean(x, axis=None, keepdims=False):
    x = _t(x)
    return _red(_torch.mean, x if x.is_floating_point() else x.to(_torch.float64), axis, keepdims)
def max(x, axis=None): return _red(_torch.amax, x, axis) if axis is not None else _torch.max(_t(x))
def min(x, axis=None): return _red(_torch.amin, x, axis) if axis is not None else _torch.min(_t(x))
def all(x, axis=None): return _torch.all(_t(x)) if axis is None else _torch.all(_t(x), dim=axis)
def any(x, axis=None): return _torch.any(_t(x)) if axis is None else _torch.any(_t(x), dim=axis)


def _pair(a, b):
    ta, tb = isinstance(a, _torch.Tensor), isinstance(b, _torch.Tensor)
    if ta and not tb:
        return a, _torch.as_tensor(b, dtype=a.dtype if isinstance(b, float) == a.is_floating_point() else None)
    if tb and not ta:
        return _torch.as_tensor(a, dtype=b.dtype if isinstance(a, float) == b.is_floating_point() else None), b
    return _t(a), _t(b)


def maximum(a, b): return _torch.maximum(*_pair(a, b))
def minimum(a, b): return _torch.minimum(*_pair(a, b))
def where(c, a, b): return _torch.where(_t(c), *_pair(a, b))
def clip(x, a_min=None, a_max=None): return _torch.clamp(_t(x), min=a_min, max=a_max)


def sort(x, axis=-1): return _torch.sort(_t(x), dim=axis).values
def quantile(x, q, axis=None): return _torch.quantile(_t(x), _t(q, _torch.float64), dim=axis, interpolation="linear")


class linalg:
    cholesky = staticmethod(lambda a: _torch.linalg.cholesky(_t(a)))
    inv = staticmethod(lambda a: _torch.linalg.inv(_t(a)))
    solve = staticmethod(lambda a, b: _torch.linalg.solve(_t(a), _t(b)))
    norm = staticmethod(lambda a, axis=None, keepdims=False: _torch.linalg.norm(_t(a), dim=axis, keepdim=keepdims))
