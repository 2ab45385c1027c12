"""Cross-chain reductions over the chain shards of a multi-GPU run.

Chains are independent, so the sampling kernels never communicate: rank r owns global
chains [r*C, (r+1)*C) (Philox counters are keyed on the global chain id, `chain_offset`).
Only the statistics the reference computes ACROSS chains cross the wire:

  mean acceptance                      tuning/adaptation.py:218, dual_averaging.py:81,199,328
  mean over chains of per-chain var    tuning/adaptation.py:238-248
  R-hat / ESS / z-score moment sums    run_benchmarks.py:161-195,108-158
  ESJD                                 samplers/GRAHMC.py:422-446 (consumer: friction tuning)

Every payload is a few doubles: one fused all-reduce (sum) per dual-averaging batch, issued
on the sampler's stream (NCCL, CUDA tensors) or on the host (gloo, CPU tensors — used by the
world_size-2 CPU tests).  Without an initialised process group every function is the identity,
so single-GPU callers run the same code.
"""
from __future__ import annotations

from typing import Sequence, Tuple

import torch
import torch.distributed as td


# collectives issued through this module since import (bench.py reports the warmup's count)
COUNTERS = {"all_reduce": 0, "all_gather": 0, "all_to_all": 0}


def is_dist() -> bool:
    return td.is_available() and td.is_initialized()


def world_size() -> int:
    return td.get_world_size() if is_dist() else 1


def rank() -> int:
    return td.get_rank() if is_dist() else 0


def chain_shard(total_chains: int) -> Tuple[int, int]:
    """(chain_offset, n_chains) of this rank: contiguous blocks, remainder to the first ranks."""
    w, r = world_size(), rank()
    base, rem = divmod(int(total_chains), w)
    n = base + (1 if r < rem else 0)
    off = r * base + min(r, rem)
    return off, n


def warm_up(device=None, n_dim: int = 20) -> None:
    """Issue every collective this package uses once, with the payload shapes of the warmup and the diagnostics, so that
    NCCL's lazy set-up (communicator, per-peer point-to-point channels of the first all-to-all, algorithm selection per
    payload size) happens here and not inside a measured or latency-sensitive call: on 8 GPUs the first all-to-all
    alone took 8 s of channel set-up in round 1 (profiles/configs_r01f_8gpu.jsonl, first config-5 line)."""
    if not (is_dist() and world_size() > 1):
        return
    w = world_size()
    dev = device if device is not None else (torch.device("cuda", torch.cuda.current_device())
                                             if td.get_backend() == "nccl" else torch.device("cpu"))
    for n in (1, 2, n_dim + 1, n_dim * 7):
        td.all_reduce(torch.zeros((n,), dtype=torch.float64, device=dev))
    td.all_reduce(torch.zeros((1,), dtype=torch.float64, device=dev), op=td.ReduceOp.MAX)
    x = torch.zeros((4, n_dim), dtype=torch.float64, device=dev)
    td.all_gather([torch.empty_like(x) for _ in range(w)], x)
    n1 = torch.zeros((1,), dtype=torch.int64, device=dev)
    td.all_gather([torch.zeros_like(n1) for _ in range(w)], n1)
    if td.get_backend() == "nccl":
        td.all_to_all([torch.empty((8,), dtype=torch.float64, device=dev) for _ in range(w)],
                      [torch.zeros((8,), dtype=torch.float64, device=dev) for _ in range(w)])
        torch.cuda.synchronize(dev)


def all_reduce_sum_(t: torch.Tensor) -> torch.Tensor:
    """In-place sum over ranks (identity without a process group)."""
    if is_dist() and world_size() > 1:
        COUNTERS["all_reduce"] += 1
        td.all_reduce(t, op=td.ReduceOp.SUM)
    return t


def all_reduce_max_(t: torch.Tensor) -> torch.Tensor:
    if is_dist() and world_size() > 1:
        COUNTERS["all_reduce"] += 1
        td.all_reduce(t, op=td.ReduceOp.MAX)
    return t


def fused_sum(parts: Sequence[torch.Tensor]) -> list:
    """Sum several small tensors over ranks with ONE all-reduce.  Everything travels as float64
    (integer counts below 2^53 stay exact).  Returns float64 tensors with the input shapes."""
    flat = [p.reshape(-1).to(torch.float64) for p in parts]
    if not (is_dist() and world_size() > 1):
        return [f.reshape(p.shape) for f, p in zip(flat, parts)]
    buf = torch.cat(flat)
    COUNTERS["all_reduce"] += 1
    td.all_reduce(buf, op=td.ReduceOp.SUM)
    out, o = [], 0
    for p in parts:
        n = p.numel()
        out.append(buf[o:o + n].reshape(p.shape))
        o += n
    return out


def global_mean(local_sum: torch.Tensor, local_count: int) -> torch.Tensor:
    """sum_ranks(local_sum) / sum_ranks(local_count); the count rides in the same payload."""
    # torch.full is a fill kernel; torch.tensor([...], device=cuda) would be a blocking H2D copy (a stream sync)
    cnt = torch.full((1,), float(local_count), dtype=torch.float64, device=local_sum.device)
    s, c = fused_sum([local_sum, cnt])
    return s / c


def mean_accept(accept_sum: torch.Tensor, n_terms: int) -> float:
    """Mean acceptance statistic over ALL chains of ALL ranks (adaptation.py:218).
    accept_sum: this shard's sum (exact integer count for HMC/GRAHMC/RWMH, fp64 sum of the
    per-transition mean accept probabilities for NUTS); n_terms: terms in that sum."""
    return float(global_mean(accept_sum.reshape(1), n_terms).item())


def mean_chain_variance(local_sum_var: torch.Tensor, local_chains: int) -> torch.Tensor:
    """(D,) mean over all chains of the per-chain variances (adaptation.py:240-248);
    local_sum_var = sum over this shard's chains (mcmc_welford_finalize)."""
    return global_mean(local_sum_var, local_chains)


def gather_chains(x: torch.Tensor, dim: int = 0, max_per_rank: int | None = None) -> torch.Tensor:
    """All-gather a (sub)sample of each rank's chains along `dim` (SW2 / rank-normalisation inputs).
    Shards may differ in size by one chain; they are padded to the max and trimmed after the gather."""
    if max_per_rank is not None and x.shape[dim] > max_per_rank:
        x = x.narrow(dim, 0, max_per_rank)
    if not (is_dist() and world_size() > 1):
        return x
    n = torch.tensor([x.shape[dim]], dtype=torch.int64, device=x.device)
    ns = [torch.zeros_like(n) for _ in range(world_size())]
    td.all_gather(ns, n)
    ns = [int(v.item()) for v in ns]
    m = max(ns)
    if x.shape[dim] < m:
        pad = list(x.shape)
        pad[dim] = m - x.shape[dim]
        x = torch.cat([x, x.new_zeros(pad)], dim=dim)
    x = x.contiguous()
    outs = [torch.empty_like(x) for _ in range(world_size())]
    td.all_gather(outs, x)
    return torch.cat([o.narrow(dim, 0, k) for o, k in zip(outs, ns)], dim=dim)


def variable_shard(n_dim: int, r: int | None = None, w: int | None = None) -> Tuple[int, int]:
    """(first variable, count) owned by rank r when the D variables are dealt in contiguous blocks."""
    w = world_size() if w is None else w
    r = rank() if r is None else r
    base, rem = divmod(int(n_dim), w)
    return r * base + min(r, rem), base + (1 if r < rem else 0)


def chains_to_variables(samples: torch.Tensor):
    """Re-shard (S, C_local, D) from chains to variables: returns ((S, C_total, d_local), [variable ids]).
    One all-to-all over NVLink (NCCL); gloo has no all-to-all, so the CPU test path all-gathers instead.
    Every rank must hold the same number of chains."""
    S, Cl, D = samples.shape
    w, r = world_size(), rank()
    d0, dn = variable_shard(D)
    if w == 1:
        return samples, list(range(D))
    if td.get_backend() == "nccl":
        # send block j = my chains of rank j's variables, (S, Cl, dn_j) contiguous
        send = [samples[:, :, variable_shard(D, j, w)[0]:sum(variable_shard(D, j, w))].contiguous() for j in range(w)]
        recv = [torch.empty((S, Cl, dn), dtype=samples.dtype, device=samples.device) for _ in range(w)]
        COUNTERS["all_to_all"] += 1
        td.all_to_all(recv, send)
    else:
        mine = samples.contiguous()
        full = [torch.empty_like(mine) for _ in range(w)]
        td.all_gather(full, mine)
        recv = [f[:, :, d0:d0 + dn] for f in full]
    return torch.cat(recv, dim=1), list(range(d0, d0 + dn))
