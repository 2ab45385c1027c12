"""A torch-fp64 stand-in for the small part of the jax API the reference's samplers and targets use.

TEST INFRASTRUCTURE ONLY (tests/golden/make_reference_golden.py).  jax 0.8.0 cannot be installed in this image
(no wheel, no network), so the reference's own source files are executed UNMODIFIED from /root/reference on top of
this module: arrays are torch.float64 CPU tensors, vmap is a Python loop over the mapped axis, value_and_grad is
torch.autograd on the per-chain scalar, lax.scan / cond / while_loop / fori_loop are plain Python control flow,
jit is the identity.  Every operation therefore runs eagerly in IEEE double precision in the order the reference's
source states it.  Random draws come from a NumPy generator keyed on the PRNG key (NOT threefry) and every draw is
appended to DRAW_LOG so the generator script can inject the same numbers into the oracle and the CUDA kernels.
"""
from __future__ import annotations

import collections
import math

import numpy as np
import torch

torch.set_default_dtype(torch.float64)  # jax_enable_x64: python floats and int/int divisions are float64

Tensor = torch.Tensor
DRAW_LOG = []   # (kind, tuple(shape), np.ndarray float64)

_DT = {float: torch.float64, int: torch.int64, bool: torch.bool, None: None}


def _dt(d):
    if isinstance(d, torch.dtype) or d is None:
        return d
    if d in _DT:
        return _DT[d]
    return {np.dtype("float64"): torch.float64, np.dtype("float32"): torch.float32, np.dtype("int32"): torch.int32,
            np.dtype("int64"): torch.int64, np.dtype("bool"): torch.bool}[np.dtype(d)]


def _t(x, dtype=None):
    if isinstance(x, Tensor):
        return x if dtype is None or x.dtype == _dt(dtype) else x.to(_dt(dtype))
    if isinstance(x, np.ndarray):
        x = torch.from_numpy(np.ascontiguousarray(x))
        return x if dtype is None else x.to(_dt(dtype))
    return torch.as_tensor(x, dtype=_dt(dtype))


if not hasattr(Tensor, "astype"):
    Tensor.astype = lambda self, d: self.to(_dt(d))


# ------------------------------------------------------------------------------------------ pytrees
def _is_namedtuple(x):
    return isinstance(x, tuple) and hasattr(x, "_fields")


def tree_flatten(tree):
    leaves = []

    def rec(t):
        if _is_namedtuple(t):
            return ("nt", type(t), [rec(c) for c in t])
        if isinstance(t, (tuple, list)):
            return ("seq", type(t), [rec(c) for c in t])
        if isinstance(t, dict):
            return ("dict", list(t.keys()), [rec(t[k]) for k in t])
        if t is None:
            return ("none",)
        leaves.append(t)
        return ("leaf",)

    return leaves, rec(tree)


def tree_unflatten(spec, leaves):
    it = iter(leaves)

    def rec(s):
        if s[0] == "nt":
            return s[1](*[rec(c) for c in s[2]])
        if s[0] == "seq":
            return s[1](rec(c) for c in s[2])
        if s[0] == "dict":
            return {k: rec(c) for k, c in zip(s[1], s[2])}
        if s[0] == "none":
            return None
        return next(it)

    return rec(spec)


def _stack_trees(trees):
    flat = [tree_flatten(t) for t in trees]
    spec = flat[0][1]
    n = len(flat[0][0])
    return tree_unflatten(spec, [torch.stack([_t(f[0][i]) for f in flat]) for i in range(n)])


# ------------------------------------------------------------------------------------------ transforms
def jit(fun=None, **_kw):
    if fun is None:
        return lambda f: f
    return fun


def vmap(fun, in_axes=0, out_axes=0):
    def mapped(*args):
        axes = in_axes if isinstance(in_axes, (tuple, list)) else (in_axes,) * len(args)
        n = next(_t(a).shape[ax] for a, ax in zip(args, axes) if ax is not None)
        outs = []
        for i in range(n):
            outs.append(fun(*[a if ax is None else _t(a).select(ax, i) for a, ax in zip(args, axes)]))
        return _stack_trees(outs)
    return mapped


def value_and_grad(fun, argnums=0, has_aux=False):
    assert argnums == 0 and not has_aux

    def vg(x, *rest):
        xr = _t(x).detach().clone().requires_grad_(True)
        with torch.enable_grad():
            y = fun(xr, *rest)
            (g,) = torch.autograd.grad(y, xr, allow_unused=True)
        if g is None:
            g = torch.zeros_like(xr)
        return y.detach(), g.detach()
    return vg


def grad(fun, argnums=0):
    return lambda x, *r: value_and_grad(fun, argnums)(x, *r)[1]


class _Config:
    def update(self, *a, **k):
        pass


config = _Config()


# ------------------------------------------------------------------------------------------ lax
class lax:
    @staticmethod
    def scan(f, init, xs=None, length=None):
        carry, ys = init, []
        if xs is None:
            seq = [None] * int(length)
        else:
            leaves, spec = tree_flatten(xs)
            n = leaves[0].shape[0]
            seq = [tree_unflatten(spec, [l[i] for l in leaves]) for i in range(n)]
        for x in seq:
            carry, y = f(carry, x)
            ys.append(y)
        if not ys or all(len(tree_flatten(y)[0]) == 0 for y in ys):
            return carry, None
        return carry, _stack_trees(ys)

    @staticmethod
    def cond(pred, true_fun, false_fun, *operands):
        return true_fun(*operands) if bool(pred) else false_fun(*operands)

    @staticmethod
    def while_loop(cond_fun, body_fun, init_val):
        val = init_val
        while bool(cond_fun(val)):
            val = body_fun(val)
        return val

    @staticmethod
    def fori_loop(lower, upper, body_fun, init_val):
        val = init_val
        for i in range(int(lower), int(upper)):
            val = body_fun(i, val)
        return val


# ------------------------------------------------------------------------------------------ random
class random:
    @staticmethod
    def PRNGKey(seed):
        return torch.tensor([0, int(seed) & 0xFFFFFFFF], dtype=torch.int64)

    key = PRNGKey

    @staticmethod
    def _ss(key, extra=()):
        k = [int(v) for v in _t(key).reshape(-1).tolist()]
        return np.random.SeedSequence(k + list(extra))

    @staticmethod
    def split(key, num=2):
        st = random._ss(key, (0x5EED,)).generate_state(2 * int(num), dtype=np.uint32).astype(np.int64)
        return torch.from_numpy(st.reshape(int(num), 2))

    @staticmethod
    def fold_in(key, data):
        st = random._ss(key, (0xF01D, int(data))).generate_state(2, dtype=np.uint32).astype(np.int64)
        return torch.from_numpy(st)

    @staticmethod
    def _gen(key):
        return np.random.Generator(np.random.Philox(random._ss(key)))

    @staticmethod
    def normal(key, shape=(), dtype=torch.float64):
        shape = tuple(int(s) for s in (shape if isinstance(shape, (tuple, list, torch.Size)) else (shape,)))
        z = random._gen(key).standard_normal(shape)
        DRAW_LOG.append(("normal", shape, np.array(z, dtype=np.float64)))
        return _t(np.asarray(z, dtype=np.float64)).to(_dt(dtype)).reshape(shape)

    @staticmethod
    def uniform(key, shape=(), dtype=torch.float64, minval=0.0, maxval=1.0):
        shape = tuple(int(s) for s in (shape if isinstance(shape, (tuple, list, torch.Size)) else (shape,)))
        u = random._gen(key).random(shape)
        DRAW_LOG.append(("uniform", shape, np.array(u, dtype=np.float64)))
        u = _t(np.asarray(u, dtype=np.float64)).reshape(shape)
        if minval != 0.0 or maxval != 1.0:
            u = minval + (maxval - minval) * u
        return u.to(_dt(dtype))

    @staticmethod
    def bernoulli(key, p=0.5, shape=None):
        # jax.random.bernoulli(key, p) is uniform(key, shape) < p
        u = random._gen(key).random(() if shape is None else tuple(shape))
        DRAW_LOG.append(("bernoulli", () if shape is None else tuple(shape), np.array(u, dtype=np.float64)))
        return _t(np.asarray(u, dtype=np.float64)) < p
