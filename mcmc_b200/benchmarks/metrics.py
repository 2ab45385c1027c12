"""Sliced Wasserstein-2 distance between MCMC draws and ground-truth draws — the reference's
benchmarks/metrics.py:12-146 on the device.

Projections are one fp64 GEMM (n x D) @ (D x P) and a device sort per projection batch; with
equal sample counts the 1-D W2 is the RMS difference of the sorted projections
(metrics.py:66-76).  Directions come from the host Philox stream (mcmc_b200.random), keyed
`fold_in(key, batch_idx)` like the reference; they are NOT JAX's threefry draws, so SW2 values
agree with the reference statistically, not bit for bit.
"""
from __future__ import annotations

from typing import Optional

import numpy as np
import torch

from .. import dist
from ..random import PRNGKey, as_key, fold_in, normal, split
from .targets import get_reference_sampler


def _dev(x) -> torch.Tensor:
    if isinstance(x, torch.Tensor):
        if not x.is_cuda:
            raise RuntimeError("mcmc_b200 metrics run on CUDA tensors only (there is no CPU fallback)")
        return x.to(torch.float64)
    if not torch.cuda.is_available():
        raise RuntimeError("mcmc_b200 metrics need a CUDA device (there is no CPU fallback)")
    return torch.from_numpy(np.ascontiguousarray(x, dtype=np.float64)).cuda()


def sliced_wasserstein_distance(samples1, samples2, n_projections: int = 500, key=None,
                                projection_batch_size: int = 100, directions=None) -> float:
    """metrics.py:12-84.  `directions` (n_projections, dim), unnormalised, replaces the drawn projection directions
    (the hook the parity test uses to feed the reference run's own directions)."""
    if key is None:
        key = PRNGKey(30)
    s1, s2 = _dev(samples1), _dev(samples2)
    s2 = s2.to(s1.device)
    n1, dim = s1.shape
    n2 = s2.shape[0]
    n_batches = (n_projections + projection_batch_size - 1) // projection_batch_size
    w2 = []
    for b in range(n_batches):
        cur = min((b + 1) * projection_batch_size, n_projections) - b * projection_batch_size
        if directions is not None:
            dirs = np.asarray(directions, dtype=np.float64)[b * projection_batch_size:b * projection_batch_size + cur]
        else:
            dirs = normal(fold_in(key, b), (cur, dim))
        dirs = dirs / np.linalg.norm(dirs, axis=1, keepdims=True)
        dt = torch.from_numpy(dirs).to(s1.device)
        p1 = torch.sort(s1 @ dt.T, dim=0).values
        p2 = torch.sort(s2 @ dt.T, dim=0).values
        if n1 != n2:
            nq = min(n1, n2)
            qs = torch.linspace(0, 1, nq, dtype=torch.float64, device=s1.device)
            p1 = _sorted_quantiles(p1, qs)
            p2 = _sorted_quantiles(p2, qs)
        w2.append(torch.sqrt(torch.mean((p1 - p2) ** 2, dim=0)))
    return float(torch.cat(w2).mean().item())


def _sorted_quantiles(p_sorted: torch.Tensor, qs: torch.Tensor) -> torch.Tensor:
    """Linear-interpolation quantiles of already sorted columns (jnp.quantile default)."""
    n = p_sorted.shape[0]
    pos = qs * (n - 1)
    lo = pos.floor().long()
    hi = torch.clamp(lo + 1, max=n - 1)
    f = (pos - lo.to(torch.float64)).unsqueeze(1)
    return p_sorted[lo] + f * (p_sorted[hi] - p_sorted[lo])


def compute_sliced_w2(samples, target_name: str, dim: int, n_reference: int = 50000, n_projections: int = 500,
                      projection_batch_size: int = 100, key=None, distributed: bool = False,
                      **target_kwargs) -> Optional[float]:
    """metrics.py:87-146: SW2 of the draws (n_samples, n_chains, dim) or (n, dim) against `n_reference`
    ground-truth draws; None if the target has no reference sampler.  With distributed=True every rank
    contributes an equal share of the subsample (all-gathered over the chain shards)."""
    if key is None:
        key = PRNGKey(123)
    ref_sampler = get_reference_sampler(target_name, dim, **target_kwargs)
    if ref_sampler is None:
        return None
    flat = _dev(samples).reshape(-1, dim)
    key, sub = split(key)
    reference = _dev(np.asarray(ref_sampler(sub, n_reference)))
    w = dist.world_size() if distributed else 1
    share = (n_reference + w - 1) // w
    key, sub = split(key)
    if flat.shape[0] > share:
        # choice without replacement (metrics.py:135), drawn ON the device: a keyed random permutation of the row indices
        # (a CPU randperm over 16.8 M rows per rank cost more than the whole distance)
        k = as_key(sub)
        g = torch.Generator(device=flat.device)
        g.manual_seed(((int(k[0]) << 32 | int(k[1])) ^ (dist.rank() if distributed else 0)) & 0x7FFFFFFFFFFFFFFF)
        idx = torch.randperm(flat.shape[0], generator=g, device=flat.device)[:share]
        flat = flat[idx]
    if w > 1:
        flat = dist.gather_chains(flat)[:n_reference]
    key, sub = split(key)
    return sliced_wasserstein_distance(flat, reference.to(flat.device), n_projections=n_projections,
                                       projection_batch_size=projection_batch_size, key=sub)
