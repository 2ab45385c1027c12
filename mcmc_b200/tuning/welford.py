"""Welford's online mean / diagonal variance, the reference's functional API
(tuning/welford.py:13-100) on torch tensors (any device, float64).

The warmup does NOT call these per sample: the per-chain Welford states of an adaptation
window live in HBM and are updated inside the sampler kernels (MCMC_FLAG_ACCUM_WELFORD,
mcmc_b200/csrc/hmc_kernel.cuh), which is what removes the reference's O(S*C) Python loop
(tuning/adaptation.py:229-235).  These functions are the API-compatible host mirror and the
specification the fused path is tested against.
"""
from __future__ import annotations

from typing import NamedTuple, Tuple

import torch


class WelfordState(NamedTuple):
    """tuning/welford.py:13-21."""
    count: torch.Tensor  # float64 scalar
    mean: torch.Tensor   # (n_dim,)
    m2: torch.Tensor     # (n_dim,)


def welford_init(n_dim: int, dtype=torch.float64, device=None) -> WelfordState:
    """tuning/welford.py:24-38."""
    return WelfordState(count=torch.zeros((), dtype=dtype, device=device),
                        mean=torch.zeros(n_dim, dtype=dtype, device=device),
                        m2=torch.zeros(n_dim, dtype=dtype, device=device))


def welford_update(state: WelfordState, x) -> WelfordState:
    """One new sample x (n_dim,) (tuning/welford.py:42-59)."""
    x = torch.as_tensor(x, dtype=state.mean.dtype, device=state.mean.device)
    count = state.count + 1.0
    delta = x - state.mean
    mean = state.mean + delta / count
    delta2 = x - mean
    m2 = state.m2 + delta * delta2
    return WelfordState(count, mean, m2)


def welford_update_batch(state: WelfordState, batch) -> WelfordState:
    """Sequential fold of a (batch_size, n_dim) batch (tuning/welford.py:63-82)."""
    batch = torch.as_tensor(batch, dtype=state.mean.dtype, device=state.mean.device)
    for i in range(batch.shape[0]):
        state = welford_update(state, batch[i])
    return state


def welford_covariance(state: WelfordState) -> Tuple[torch.Tensor, torch.Tensor]:
    """(mean, variance) with variance = m2 / (max(count, 2) - 1) (tuning/welford.py:86-100)."""
    n = torch.clamp(state.count, min=2.0)
    return state.mean, state.m2 / (n - 1.0)
