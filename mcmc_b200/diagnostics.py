"""Convergence diagnostics and quality gates on the device.

Mirrors the reference's `compute_diagnostics` / `check_summary_statistics` / `detect_divergences`
(run_benchmarks.py:161-195, 108-158, 90-103) and the gate thresholds (:55-62, :777-804).  The
reference delegates the statistics to ArviZ (third party); here they are the published algorithm
(Vehtari et al. 2021) run by the kernels of mcmc_b200/csrc/diag.cu:

  per variable:  sort (device radix sort) -> rank_normalize kernel (average ranks over ties,
                 z = Phi^-1((r-3/8)/(N+1/4))) -> split_chain_stats kernel (per-sequence mean +
                 autocovariance at every lag, summed over sequences) -> R-hat / Geyer ESS from the
                 (n_lags + 3) reduced doubles on the host.

Multi-GPU: chains are sharded, but rank-normalisation needs the ranks of a variable over ALL
chains.  `compute_diagnostics(..., distributed=True)` therefore re-shards from chains to
variables with one all-to-all (each rank receives every chain of its D/W variables), computes
those variables completely, and all-gathers the D results (mcmc_b200/dist.py).
"""
from __future__ import annotations

import ctypes as C
import math
from typing import Dict, Optional

import numpy as np
import torch

from . import _abi, dist

# gate thresholds, run_benchmarks.py:55-62
MIN_ESS_HARD_GATE = 400
MIN_ESS_TAIL_HARD_GATE = 100
MIN_ESS_QUALITY = 400
MIN_ESS_TAIL_QUALITY = 200
INEFFICIENT_THRESHOLD = 0.01
HIGH_EFFICIENCY_THRESHOLD = 0.1


def _ptr(t):
    return None if t is None else C.c_void_p(t.data_ptr())


def _stream(dev):
    return C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)


def _require_cuda(x: torch.Tensor):
    if not (isinstance(x, torch.Tensor) and x.is_cuda):
        raise RuntimeError("mcmc_b200.diagnostics runs on CUDA tensors only (there is no CPU fallback)")


def rank_normalize(x: torch.Tensor) -> torch.Tensor:
    """ArviZ `_z_scale`: rank-normalise ALL elements of x jointly (average ranks for ties)."""
    _require_cuda(x)
    L = _abi.lib()
    flat = x.reshape(-1).contiguous()
    vals, perm = torch.sort(flat)
    out = torch.empty_like(flat)
    _abi.check(L.mcmc_set_device(x.device.index))
    _abi.check(L.mcmc_rank_normalize(_ptr(vals), _ptr(perm), flat.numel(), _ptr(out), _stream(x.device)))
    return out.reshape(x.shape), vals


def rank_normalize_fold(x: torch.Tensor):
    """ArviZ `_z_scale(x)` and `_z_scale(abs(x - median(x)))` over ALL elements of x from one sort
    (mcmc_rank_normalize_fold).  Returns (z_bulk, z_fold, sorted values); bit-identical to two rank_normalize calls."""
    _require_cuda(x)
    L = _abi.lib()
    flat = x.reshape(-1).contiguous()
    n = flat.numel()
    if n >= 2 ** 31:
        z, vals = rank_normalize(x)
        med = 0.5 * (vals[(n - 1) // 2] + vals[n // 2])
        return z, rank_normalize((x - med).abs())[0], vals
    nb = C.c_size_t(0)
    _abi.check(L.mcmc_rank_normalize_fold_workspace(n, C.byref(nb)))
    ws = torch.empty((nb.value,), dtype=torch.uint8, device=x.device)
    vals, zb, zf = torch.empty_like(flat), torch.empty_like(flat), torch.empty_like(flat)
    _abi.check(L.mcmc_set_device(x.device.index))
    _abi.check(L.mcmc_rank_normalize_fold(_ptr(flat), n, _ptr(vals), _ptr(zb), _ptr(zf), _ptr(ws), nb.value, _stream(x.device)))
    ws.record_stream(torch.cuda.current_stream(x.device))
    return zb.reshape(x.shape), zf.reshape(x.shape), vals


def split_chain_stats(y: torch.Tensor, half_len: int, second_half_row: int, n_lags: Optional[int] = None):
    """y: (rows, C) float64 CUDA tensor (any strides).  Returns (acov_sum (n_lags,), mean_stats (3,)) on the device."""
    _require_cuda(y)
    L = _abi.lib()
    rows, Cn = y.shape
    n_lags = int(half_len if n_lags is None else min(n_lags, half_len))
    acov = torch.empty((n_lags,), dtype=torch.float64, device=y.device)
    stats = torch.empty((3,), dtype=torch.float64, device=y.device)
    nb = C.c_size_t(0)
    _abi.check(L.mcmc_split_chain_stats_workspace(int(half_len), int(Cn), n_lags, C.byref(nb)))
    ws = torch.empty((nb.value,), dtype=torch.uint8, device=y.device)
    _abi.check(L.mcmc_set_device(y.device.index))
    _abi.check(L.mcmc_split_chain_stats(_ptr(y), int(half_len), int(second_half_row), int(Cn), int(y.stride(0)),
                                        int(y.stride(1)), n_lags, _ptr(acov), _ptr(stats), _ptr(ws), nb.value,
                                        _stream(y.device)))
    ws.record_stream(torch.cuda.current_stream(y.device))
    return acov, stats


# ---------------------------------------------------------------------------------------------
# host-side finish on the reduced sums (ArviZ `_rhat`, `_ess`)

def _rhat_from(acov0_sum: float, stats, n_draw: int) -> float:
    cnt, _, m2 = stats
    within = (acov0_sum / cnt) * n_draw / (n_draw - 1.0)            # mean of the ddof=1 chain variances
    between = n_draw * (m2 / (cnt - 1.0))
    return math.sqrt((between / within + n_draw - 1.0) / n_draw)


def _ess_from(acov_sum: np.ndarray, stats, n_draw: int) -> float:
    """Geyer initial positive + monotone sequence on the chain-averaged autocovariance."""
    cnt, _, m2 = stats
    n_chain = int(cnt)
    if acov_sum[0] <= 0.0:                                          # constant input: ArviZ returns the size
        return float(n_chain * n_draw)
    acov = acov_sum / cnt
    n_avail = acov.shape[0]
    mean_var = acov[0] * n_draw / (n_draw - 1.0)
    var_plus = mean_var * (n_draw - 1.0) / n_draw
    if n_chain > 1:
        var_plus += m2 / (cnt - 1.0)
    rho = np.zeros(n_draw)
    rho_even = 1.0
    rho[0] = rho_even
    rho_odd = 1.0 - (mean_var - acov[1]) / var_plus
    rho[1] = rho_odd
    t = 1
    while t < (n_draw - 3) and (rho_even + rho_odd) > 0.0:
        if t + 2 >= n_avail:
            raise RuntimeError(f"autocovariance truncated at lag {n_avail} before Geyer's sequence ended; "
                               "pass a larger max_lag")
        rho_even = 1.0 - (mean_var - acov[t + 1]) / var_plus
        rho_odd = 1.0 - (mean_var - acov[t + 2]) / var_plus
        if (rho_even + rho_odd) >= 0:
            rho[t + 1] = rho_even
            rho[t + 2] = rho_odd
        t += 2
    max_t = t - 2
    if rho_even > 0:
        rho[max_t + 1] = rho_even
    t = 1
    while t <= max_t - 2:
        if (rho[t + 1] + rho[t + 2]) > (rho[t - 1] + rho[t]):
            rho[t + 1] = (rho[t - 1] + rho[t]) / 2.0
            rho[t + 2] = rho[t + 1]
        t += 2
    ess = n_chain * n_draw
    tau = -1.0 + 2.0 * float(np.sum(rho[: max_t + 1])) + float(np.sum(rho[max_t + 1: max_t + 2]))
    tau = max(tau, 1.0 / math.log10(ess))
    return float("nan") if np.isnan(rho).any() else float(ess / tau)


def _split_rows(col: torch.Tensor):
    """(S, C) -> contiguous (2n, C): first n rows and last n rows (ArviZ `_split_chains`)."""
    S = col.shape[0]
    n = S // 2
    if S == 2 * n:
        return col.contiguous(), n
    return torch.cat([col[:n], col[S - n:]], dim=0).contiguous(), n


def _variable_device(col: torch.Tensor, max_lag: Optional[int] = None):
    """Device phase of one variable: enqueues the sorts and kernels and returns (packed device tensor, n_lags, n) without
    waiting for the GPU.  Layout of the packed tensor: see _variable_finish."""
    _require_cuda(col)
    S, Cn = col.shape
    if S < 4:
        raise ValueError("need at least 4 draws per chain")
    split, n = _split_rows(col.to(torch.float64))
    # bulk: rank-normalised split chains; tail R-hat: folded around the median of the split draws — one sort for both
    z, zf, sorted_vals = rank_normalize_fold(split)
    a_bulk, s_bulk = split_chain_stats(z, n, n, max_lag)
    N = sorted_vals.numel()
    a_fold, s_fold = split_chain_stats(zf, n, n, 1)
    # tail ESS: indicators of the 5% / 95% quantiles of ALL draws
    if S == 2 * n:
        lo = (int((N - 1) * 0.05), int((N - 1) * 0.95))                      # floor of q*(N-1): host integers, no sync
        fr = ((N - 1) * 0.05 - lo[0], (N - 1) * 0.95 - lo[1])
        q = [sorted_vals[lo[k]] + fr[k] * (sorted_vals[min(lo[k] + 1, N - 1)] - sorted_vals[lo[k]]) for k in range(2)]   # np.quantile 'linear'
    else:
        qq = torch.quantile(col.reshape(-1).to(torch.float64), torch.tensor([0.05, 0.95], dtype=torch.float64).to(col.device, non_blocking=True))
        q = [qq[0], qq[1]]
    tails = []
    for k in range(2):
        ind = (split <= q[k]).to(torch.float64)
        tails.append(split_chain_stats(ind, n, n, max_lag))
    # mean ESS / mcse on the raw split chains
    a_raw, s_raw = split_chain_stats(split, n, n, max_lag)
    flat = col.reshape(-1).to(torch.float64)
    mean = flat.mean()
    sd = flat.std(unbiased=True)
    packed = torch.cat([a_bulk, s_bulk, a_fold[:1], s_fold, tails[0][0], tails[0][1], tails[1][0], tails[1][1], a_raw,
                        s_raw, mean.reshape(1), sd.reshape(1)])
    return packed, a_bulk.numel(), n


def _variable_finish(host: np.ndarray, nl: int, n: int) -> Dict[str, float]:
    """Host phase: R-hat and Geyer ESS from the reduced sums of one variable (ArviZ `_rhat`, `_ess`)."""
    o = 0

    def take(k):
        nonlocal o
        v = host[o:o + k]
        o += k
        return v
    ab, sb = take(nl), take(3)
    af, sf = take(1), take(3)
    at0, st0 = take(nl), take(3)
    at1, st1 = take(nl), take(3)
    ar, sr = take(nl), take(3)
    mean_h, sd_h = float(take(1)[0]), float(take(1)[0])
    rhat = max(_rhat_from(ab[0], sb, n), _rhat_from(af[0], sf, n))
    ess_mean = _ess_from(ar, sr, n)
    return {"rhat": rhat, "ess_bulk": _ess_from(ab, sb, n), "ess_tail": min(_ess_from(at0, st0, n), _ess_from(at1, st1, n)),
            "ess_mean": ess_mean, "mean": mean_h, "sd": sd_h, "mcse_mean": sd_h / math.sqrt(ess_mean)}


def variable_diagnostics(col: torch.Tensor, max_lag: Optional[int] = None) -> Dict[str, float]:
    """All statistics of ONE variable from its (S, C_all) draws: rank-normalised split R-hat,
    bulk / tail / mean ESS, mean, sd (ddof=1), mcse_mean."""
    packed, nl, n = _variable_device(col, max_lag)
    return _variable_finish(packed.cpu().numpy(), nl, n)


_KEYS = ("rhat", "ess_bulk", "ess_tail", "ess_mean", "mean", "sd", "mcse_mean")


def compute_diagnostics(samples: torch.Tensor, distributed: bool = False, max_lag: Optional[int] = None) -> Dict:
    """samples (n_samples, n_chains, n_dim) CUDA float64 -> the dict of run_benchmarks.py:161-195.

    `summary` holds arrays `mean`, `sd`, `mcse_mean` (unrounded; az.summary's default display rounding to 3
    decimals is not applied — see check_summary_statistics).  With distributed=True `samples` is this rank's
    chain shard and every rank returns the statistics over ALL chains."""
    _require_cuda(samples)
    S, Cn, D = samples.shape
    if distributed and dist.world_size() > 1:
        cols, dims = dist.chains_to_variables(samples)
    else:
        cols, dims = samples, list(range(D))
    local = torch.zeros((D, len(_KEYS)), dtype=torch.float64)
    # every variable's device work is enqueued before the first result is read: ONE wait for the GPU, and the host's
    # Geyer loops run after it instead of stalling the stream twenty times
    pending = [_variable_device(cols[:, :, j], max_lag) for j in range(len(dims))]
    if pending:
        host = torch.stack([p[0] for p in pending]).cpu().numpy()
        for j, d in enumerate(dims):
            r = _variable_finish(host[j], pending[j][1], pending[j][2])
            local[d] = torch.tensor([r[k] for k in _KEYS], dtype=torch.float64)
    if distributed and dist.world_size() > 1:
        t = local.to(samples.device)
        dist.all_reduce_sum_(t)          # every variable is filled by exactly one rank, zeros elsewhere
        local = t.cpu()
    tab = {k: local[:, i].numpy().copy() for i, k in enumerate(_KEYS)}
    n_all = S * cols.shape[1]
    pop_sd = tab["sd"] * math.sqrt((n_all - 1) / n_all)      # np.std (ddof=0), run_benchmarks.py:193
    return {
        "rhat_max": float(np.max(tab["rhat"])), "rhat_mean": float(np.mean(tab["rhat"])),
        "ess_bulk_min": float(np.min(tab["ess_bulk"])), "ess_bulk_mean": float(np.mean(tab["ess_bulk"])),
        "ess_tail_min": float(np.min(tab["ess_tail"])), "ess_tail_mean": float(np.mean(tab["ess_tail"])),
        "mean_estimate": tab["mean"], "std_estimate": pop_sd,
        "rhat": tab["rhat"], "ess_bulk": tab["ess_bulk"], "ess_tail": tab["ess_tail"],
        "summary": {"mean": tab["mean"], "sd": tab["sd"], "mcse_mean": tab["mcse_mean"]},
    }


def detect_divergences(delta_H: torch.Tensor, threshold: float = 1000.0) -> torch.Tensor:
    """run_benchmarks.py:90-103."""
    return delta_H.abs() > threshold


def _norm_ppf(p: float) -> float:
    """Phi^-1 by Newton on erfc (host scalar; scipy.stats.norm.ppf at run_benchmarks.py:142)."""
    x = 0.0
    for _ in range(60):
        cdf = 0.5 * math.erfc(-x / math.sqrt(2.0))
        pdf = math.exp(-0.5 * x * x) / math.sqrt(2.0 * math.pi)
        step = (cdf - p) / pdf
        x -= step
        if abs(step) < 1e-15:
            break
    return x


def check_summary_statistics(diagnostics: Dict, target, significance: float = 0.05, arviz_rounding: bool = False) -> Dict:
    """Bonferroni z-test of the estimated means against target.true_mean (run_benchmarks.py:108-158).

    arviz_rounding=True reproduces the reference's inputs literally: it reads `mean` and `mcse_mean` out of
    az.summary's DataFrame, which is rounded to 3 decimals — with many chains mcse_mean rounds to 0.000 and the
    z-score degenerates, so the default uses the unrounded values."""
    if target.true_mean is None or target.true_cov is None:
        return {"pass": True, "max_z": 0.0, "threshold": None, "reason": "No ground truth"}
    means = np.asarray(diagnostics["summary"]["mean"], dtype=np.float64)
    mcse = np.asarray(diagnostics["summary"]["mcse_mean"], dtype=np.float64)
    if arviz_rounding:
        means, mcse = np.round(means, 3), np.round(mcse, 3)
    true_mean = np.asarray(target.true_mean, dtype=np.float64)
    n_dim = len(means)
    z_threshold = _norm_ppf(1 - (significance / n_dim) / 2)
    epsilon = 1e-8 * np.maximum(np.abs(true_mean), 1.0) + 1e-16
    z_scores = (means - true_mean) / (mcse + epsilon)
    max_z = float(np.max(np.abs(z_scores)))
    passed = max_z < z_threshold
    return {"pass": passed, "max_z": max_z, "threshold": z_threshold, "z_scores": z_scores,
            "reason": None if passed else f"max |z|={max_z:.2f} > {z_threshold:.2f}"}


def quality_gates(diagnostics: Dict, num_samples: int, divergence_rate: float = 0.0, stats_pass: bool = True,
                  has_true_mean: bool = True) -> Dict:
    """The usable / quality gates of run_benchmarks.py:777-804."""
    rhat_max, ess_min, ess_tail_min = diagnostics["rhat_max"], diagnostics["ess_bulk_min"], diagnostics["ess_tail_min"]
    ess_per_sample = ess_min / num_samples
    usable = (rhat_max < 1.05 and ess_min >= MIN_ESS_HARD_GATE and ess_tail_min >= MIN_ESS_TAIL_HARD_GATE
              and divergence_rate < 0.05)
    if usable:
        quality = (rhat_max < 1.01 and ess_min >= MIN_ESS_QUALITY and ess_tail_min >= MIN_ESS_TAIL_QUALITY
                   and divergence_rate < 0.01 and (not has_true_mean or stats_pass))
        ineff, high = ess_per_sample < INEFFICIENT_THRESHOLD, ess_per_sample > HIGH_EFFICIENCY_THRESHOLD
    else:
        quality, ineff, high = False, False, False
    return {"usable": bool(usable), "quality_pass": bool(quality), "overall_pass": bool(quality),
            "is_inefficient": bool(ineff), "is_high_efficiency": bool(high), "ess_per_sample": ess_per_sample}
