"""The native random stream on the device (mcmc_philox_fill = the draws the sampler kernels make) against an
independent evaluation: Philox4x32-10 in NumPy with the counter layout of oracle/rng_common.h, the Box-Muller
transform in 80-bit long double.  Pins the table-driven sin/cos, the one-step square root and the integer-argument
log of mcmc_b200/csrc/fastmath.cuh ON THE GPU (the MUFU seed differs from the host build's), and moments at size."""
import numpy as np
import pytest
import torch

from mcmc_b200.random import philox4x32_10

pytestmark = pytest.mark.gpu


def _reference(key, C, D, it0, n_it, off):
    gid = (np.arange(C, dtype=np.uint64) + np.uint64(off))
    lo, hi = (gid & np.uint64(0xFFFFFFFF)).astype(np.uint32), (gid >> np.uint64(32)).astype(np.uint32)
    z = np.zeros((n_it, C, D), dtype=np.longdouble)
    lu = np.zeros((n_it, C), dtype=np.longdouble)
    u53 = lambda h, l: (h.astype(np.uint64) << np.uint64(21)) | (l.astype(np.uint64) >> np.uint64(11))
    two_pi = 2 * np.longdouble("3.14159265358979323846264338327950288")
    for t in range(n_it):
        itv = np.full(C, it0 + t, dtype=np.uint32)
        for j in range((D + 1) // 2):
            o = philox4x32_10(lo, hi, itv, np.full(C, j, dtype=np.uint32), key[0], key[1])
            u1 = (u53(o[0], o[1]) + np.uint64(1)).astype(np.longdouble) / np.longdouble(2.0 ** 53)
            u2 = u53(o[2], o[3]).astype(np.longdouble) / np.longdouble(2.0 ** 53)
            rad = np.sqrt(-2 * np.log(u1))
            z[t, :, 2 * j] = rad * np.cos(two_pi * u2)
            if 2 * j + 1 < D:
                z[t, :, 2 * j + 1] = rad * np.sin(two_pi * u2)
        o = philox4x32_10(lo, hi, itv, np.full(C, 1 << 24, dtype=np.uint32), key[0], key[1])
        with np.errstate(divide="ignore"):
            lu[t] = np.log(u53(o[0], o[1]).astype(np.longdouble) / np.longdouble(2.0 ** 53))
    return z, lu


@pytest.mark.parametrize("D,off", [(20, 0), (7, (1 << 32) - 100)])
def test_device_stream_matches_long_double_box_muller(gpu_lib, D, off):
    from mcmc_b200.runtime import philox_fill
    key = np.array([0xDEADBEEF, 12345], dtype=np.uint32)
    C, it0, n_it = 3000, 5, 3
    z, lu = philox_fill(key, C, D, it0, n_it, chain_offset=off)
    zr, lur = _reference(key, C, D, it0, n_it, off)
    z, lu = z.cpu().numpy(), lu.cpu().numpy()
    # |z| = rad * |cos|: the transform's error is relative in rad and ABSOLUTE in the angle part (<= 1.7e-16)
    rad = np.sqrt(zr[..., 0::2][..., : D // 2] ** 2 + zr[..., 1::2] ** 2) if D % 2 == 0 else None
    err = np.abs(z.astype(np.longdouble) - zr)
    assert float(err.max()) <= 2.5e-15, float(err.max())             # rad <= 8.6
    if rad is not None:
        scale = np.repeat(np.asarray(rad, dtype=np.float64), 2, axis=-1)
        assert float((err / (1e-300 + scale)).max()) <= 8e-16
    rel = np.abs(lu.astype(np.longdouble) - lur) / np.maximum(np.abs(lur), np.longdouble(1e-12))
    assert float(rel.max()) <= 4.5e-16


def test_device_stream_moments_at_size(gpu_lib):
    """2.6e8 draws: mean, variance, skewness, excess kurtosis and the 4.5-sigma tail mass of the normals; mean of the
    log-uniform (E log U = -1); lag-1 correlation across dimensions and across iterations."""
    from mcmc_b200.runtime import philox_fill
    C, D, n_it = 1 << 20, 20, 12
    z, lu = philox_fill(7, C, D, 0, n_it)
    n = z.numel()
    m = float(z.mean())
    v = float(z.var())
    sk = float((z ** 3).mean())
    ku = float((z ** 4).mean()) - 3.0
    assert abs(m) < 5 / n ** 0.5 and abs(v - 1) < 5 * (2 / n) ** 0.5
    assert abs(sk) < 5 * (15 / n) ** 0.5 and abs(ku) < 5 * (96 / n) ** 0.5
    tail = float((z.abs() > 4.5).double().mean())
    p = 6.795346249477e-06                                             # 2 * (1 - Phi(4.5))
    assert abs(tail - p) < 5 * (p / n) ** 0.5
    assert abs(float(lu.mean()) + 1.0) < 5 / lu.numel() ** 0.5
    c_dim = float((z[..., :-1] * z[..., 1:]).mean())
    c_it = float((z[:-1] * z[1:]).mean())
    assert abs(c_dim) < 5 / n ** 0.5 and abs(c_it) < 5 / n ** 0.5
