"""Host-side PRNG keys (the role jax.random plays for the reference's callers).

A key is a uint32[2] array, like JAX's threefry key; it is used verbatim as the
Philox4x32-10 key of the sampling kernels (mcmc_b200/csrc/philox.cuh), and
`split` / `fold_in` derive child keys with the same Philox function, so every
stream stays counter-based and reproducible.  `normal` / `uniform` / `gamma`
serve the small host-side draws of the reference's init samplers
(benchmarks/targets.py:187-193,253-255,313-316,389-393,456-465).
"""
from __future__ import annotations

import numpy as np

_M0, _M1 = np.uint64(0xD2511F53), np.uint64(0xCD9E8D57)
_W0, _W1 = np.uint32(0x9E3779B9), np.uint32(0xBB67AE85)
_TAG_SPLIT, _TAG_DRAW = np.uint32(0xF0000000), np.uint32(0xE0000000)


def philox4x32_10(c0, c1, c2, c3, k0, k1):
    """Vectorised Philox4x32-10 over uint32 arrays (Salmon et al., SC'11)."""
    c0, c1, c2, c3 = (np.asarray(x, dtype=np.uint32) for x in (c0, c1, c2, c3))
    k0 = np.uint32(k0)
    k1 = np.uint32(k1)
    with np.errstate(over="ignore"):
        for _ in range(10):
            p0 = _M0 * c0.astype(np.uint64)
            p1 = _M1 * c2.astype(np.uint64)
            n0 = (p1 >> np.uint64(32)).astype(np.uint32) ^ c1 ^ k0
            n2 = (p0 >> np.uint64(32)).astype(np.uint32) ^ c3 ^ k1
            c0, c1, c2, c3 = n0, p1.astype(np.uint32), n2, p0.astype(np.uint32)
            k0 = np.uint32((int(k0) + int(_W0)) & 0xFFFFFFFF)
            k1 = np.uint32((int(k1) + int(_W1)) & 0xFFFFFFFF)
    return c0, c1, c2, c3


def as_key(key) -> np.ndarray:
    """Accept an int seed, a uint32[2] array (JAX-style key) or a torch tensor."""
    if hasattr(key, "detach"):
        key = key.detach().cpu().numpy()
    if isinstance(key, (int, np.integer)):
        v = int(key) & 0xFFFFFFFFFFFFFFFF
        return np.array([v >> 32, v & 0xFFFFFFFF], dtype=np.uint32)
    k = np.asarray(key)
    if k.shape != (2,):
        raise ValueError(f"a PRNG key must be an int seed or a uint32[2] array, got shape {k.shape}")
    return k.astype(np.uint32)


def PRNGKey(seed: int) -> np.ndarray:
    return as_key(int(seed))


def split(key, num: int = 2) -> np.ndarray:
    """`num` child keys, shape (num, 2)."""
    k = as_key(key)
    i = np.arange(num, dtype=np.uint32)
    z = np.zeros(num, dtype=np.uint32)
    o = philox4x32_10(i, z, z, np.full(num, _TAG_SPLIT, dtype=np.uint32), k[0], k[1])
    return np.stack([o[0], o[1]], axis=1)


def fold_in(key, data: int) -> np.ndarray:
    k = as_key(key)
    d = np.array([int(data) & 0xFFFFFFFF], dtype=np.uint32)
    o = philox4x32_10(d, np.array([1], dtype=np.uint32), np.zeros(1, np.uint32),
                      np.array([_TAG_SPLIT], dtype=np.uint32), k[0], k[1])
    return np.array([o[0][0], o[1][0]], dtype=np.uint32)


def _blocks(key, n):
    k = as_key(key)
    i = np.arange(n, dtype=np.uint64)
    lo = (i & np.uint64(0xFFFFFFFF)).astype(np.uint32)
    hi = (i >> np.uint64(32)).astype(np.uint32)
    z = np.zeros(n, dtype=np.uint32)
    return philox4x32_10(lo, hi, z, np.full(n, _TAG_DRAW, dtype=np.uint32), k[0], k[1])


def _u53(hi, lo):
    return (hi.astype(np.uint64) << np.uint64(21)) | (lo.astype(np.uint64) >> np.uint64(11))


def uniform(key, shape=()) -> np.ndarray:
    """U[0,1) float64."""
    n = int(np.prod(shape)) if shape != () else 1
    o = _blocks(key, (n + 1) // 2)
    u = np.concatenate([_u53(o[0], o[1]), _u53(o[2], o[3])]).astype(np.float64) * 2.0 ** -53
    return u[:n].reshape(shape)


def normal(key, shape=()) -> np.ndarray:
    """N(0,1) float64 (Box-Muller on the Philox blocks, as in the kernels)."""
    n = int(np.prod(shape)) if shape != () else 1
    o = _blocks(key, (n + 1) // 2)
    u1 = (_u53(o[0], o[1]) + np.uint64(1)).astype(np.float64) * 2.0 ** -53
    u2 = _u53(o[2], o[3]).astype(np.float64) * 2.0 ** -53
    r = np.sqrt(-2.0 * np.log(u1))
    z = np.stack([r * np.cos(2.0 * np.pi * u2), r * np.sin(2.0 * np.pi * u2)], axis=1).reshape(-1)
    return z[:n].reshape(shape)


def gamma(key, a: float, shape=()) -> np.ndarray:
    """Gamma(a, 1) draws; NumPy's generator seeded from the key (host-side init only)."""
    k = as_key(key)
    g = np.random.Generator(np.random.Philox(key=int(k[0]) << 32 | int(k[1])))
    return g.gamma(a, size=shape)


def bernoulli(key, p: float = 0.5, shape=()) -> np.ndarray:
    return uniform(key, shape) < p
