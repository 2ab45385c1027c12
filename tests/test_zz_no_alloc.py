"""mcmc_welford_finalize_noalloc (the allocation-free variant whose block partials live in a file-scope `static
__device__` array of the library) against mcmc_welford_finalize (caller workspace), both through the C ABI with raw
device pointers.  Same kernels in the same order, so the results must be BIT-IDENTICAL (the tolerance of the other
welford tests, 1e-10, is asserted as well).  Two calls in a row share the static array and it keeps the first call's
contents: the second call uses a smaller n_dim and another chain count, so stale partials would show.

The file sorts last on purpose (a failure here hides nothing).  It needs no other test to
have run: every input is made here, the library is loaded by the `gpu_lib` fixture.
"""
import ctypes as C

import numpy as np
import pytest
import torch

from helpers import relerr
from mcmc_b200 import _abi

pytestmark = pytest.mark.gpu

NOALLOC_MAX_DIM = 128          # MCMC_NOALLOC_MAX_DIM of include/mcmc_b200.h


def _ptr(t):
    return C.c_void_p(t.data_ptr())


def _stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def _both(L, m2, count):
    Cn, D = m2.shape
    ws = torch.empty((256 * D * 8,), dtype=torch.uint8, device="cuda")
    want = torch.full((D,), float("nan"), dtype=torch.float64, device="cuda")
    got = torch.full((D,), float("nan"), dtype=torch.float64, device="cuda")
    _abi.check(L.mcmc_set_device(torch.cuda.current_device()))
    _abi.check(L.mcmc_welford_finalize(_ptr(m2), float(count), Cn, D, _ptr(want), _ptr(ws), ws.numel(), _stream()))
    _abi.check(L.mcmc_welford_finalize_noalloc(_ptr(m2), float(count), Cn, D, _ptr(got), _stream()))
    torch.cuda.synchronize()
    return got.cpu().numpy(), want.cpu().numpy()


def test_noalloc_entry_matches_workspace_entry_on_two_calls_in_a_row(gpu_lib):
    rng = np.random.default_rng(11)
    # call 1: wide (the maximum n_dim the static scratch holds), many chains
    m2_a = torch.from_numpy(rng.random((5000, NOALLOC_MAX_DIM)) * 40.0).cuda()
    got, want = _both(gpu_lib, m2_a, 37.0)
    assert np.isfinite(want).all()
    assert relerr(got, want) <= 1e-10
    assert np.array_equal(got, want)
    host = (m2_a.cpu().numpy() / 36.0).sum(axis=0)                # sum_c m2 / (max(count, 2) - 1)
    assert relerr(got, host) <= 1e-12
    # call 2, right after: the config-3 shape (32,768 chains x 20 dims); the static array still holds call 1's
    # 256 x 128 partials, none of which may leak into this result
    m2_b = torch.from_numpy(rng.random((32768, 20)) * 3.0).cuda()
    got, want = _both(gpu_lib, m2_b, 1.0)                         # count < 2 -> divisor max(count, 2) - 1 = 1
    assert relerr(got, want) <= 1e-10
    assert np.array_equal(got, want)
    assert relerr(got, m2_b.cpu().numpy().sum(axis=0)) <= 1e-12
    # call 3: fewer chains than reduction blocks (most blocks write 0.0 over what call 2 left)
    m2_c = torch.from_numpy(rng.random((7, 3))).cuda()
    got, want = _both(gpu_lib, m2_c, 12.0)
    assert np.array_equal(got, want)


def test_noalloc_entry_rejects_sizes_beyond_the_static_scratch(gpu_lib):
    m2 = torch.zeros((4, NOALLOC_MAX_DIM + 1), dtype=torch.float64, device="cuda")
    out = torch.zeros((NOALLOC_MAX_DIM + 1,), dtype=torch.float64, device="cuda")
    rc = gpu_lib.mcmc_welford_finalize_noalloc(_ptr(m2), 5.0, 4, NOALLOC_MAX_DIM + 1, _ptr(out), _stream())
    assert rc == _abi.ENOTSUP
    msg = gpu_lib.mcmc_last_error().decode()
    assert "exceeds the static scratch" in msg and str(NOALLOC_MAX_DIM) in msg
    assert gpu_lib.mcmc_welford_finalize_noalloc(None, 5.0, 4, 3, _ptr(out), _stream()) == _abi.EINVAL


def test_noalloc_entry_allocates_no_device_memory(gpu_lib):
    """cudaMemGetInfo before and after a burst of calls: the free memory of the device does not move (the first call's
    set-up only resolves the symbol's address)."""
    m2 = torch.ones((1024, 20), dtype=torch.float64, device="cuda")
    out = torch.empty((20,), dtype=torch.float64, device="cuda")
    _abi.check(gpu_lib.mcmc_welford_finalize_noalloc(_ptr(m2), 9.0, 1024, 20, _ptr(out), _stream()))
    torch.cuda.synchronize()
    free0, _ = torch.cuda.mem_get_info()
    for _ in range(50):
        _abi.check(gpu_lib.mcmc_welford_finalize_noalloc(_ptr(m2), 9.0, 1024, 20, _ptr(out), _stream()))
    torch.cuda.synchronize()
    free1, _ = torch.cuda.mem_get_info()
    assert free1 == free0
    assert torch.equal(out, torch.full((20,), 128.0, dtype=torch.float64, device="cuda"))   # 1024 / 8
