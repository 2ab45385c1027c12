"""NumPy/SciPy restatement of the convergence diagnostics the reference obtains from ArviZ.

TEST INFRASTRUCTURE (see oracle/oracle.py): the checker of mcmc_b200/diagnostics.py, never the
product path.

Reference call sites: run_benchmarks.py:161-195 (`compute_diagnostics`: az.rhat, az.ess bulk /
tail, az.summary mean / mcse_mean / sd) and :108-158 (`check_summary_statistics`).  ArviZ itself
(pinned arviz==0.22.0, requirements.txt:1) is a third-party dependency absent from
/root/reference and from this image, so its published algorithm is restated here (Vehtari,
Gelman, Simpson, Carpenter, Buerkner 2021, "Rank-normalization, folding, and localization"):
split chains, rank-normalise with z = Phi^-1((r - 3/8)/(N + 1/4)) (average ranks for ties), FFT
autocovariance normalised by n, Geyer initial positive + monotone sequence, tail ESS = min of
the ESS of the 5% and 95% quantile indicators, mcse_mean = sd / sqrt(ESS_mean).
PARITY UNPINNED against ArviZ itself (not installable here); pinned by known-answer cases in
tests/test_diagnostics_cpu.py (iid draws -> ESS ~ N, AR(1) -> N(1-rho)/(1+rho), R-hat -> 1).

Arrays are (n_chains, n_draws) like ArviZ's internal layout.
"""
from __future__ import annotations

import numpy as np
from scipy import fft as sfft
from scipy import stats


def split_chains(ary):
    ary = np.atleast_2d(np.asarray(ary))
    half = ary.shape[1] // 2
    return np.vstack((ary[:, :half], ary[:, -half:]))


def z_scale(ary):
    ary = np.asarray(ary)
    rank = stats.rankdata(ary, method="average")
    rank = (rank - 3.0 / 8.0) / (ary.size - 2.0 * 3.0 / 8.0 + 1.0)
    return stats.norm.ppf(rank).reshape(ary.shape)


def autocov(ary):
    n = ary.shape[1]
    m = sfft.next_fast_len(2 * n)
    ary = ary - ary.mean(axis=1, keepdims=True)
    f = np.fft.rfft(ary, n=m, axis=1)
    f *= np.conjugate(f)
    return np.fft.irfft(f, n=m, axis=1)[:, :n] / n


def rhat_plain(ary):
    ary = np.asarray(ary, dtype=float)
    n = ary.shape[1]
    chain_mean = ary.mean(axis=1)
    chain_var = ary.var(axis=1, ddof=1)
    between = n * chain_mean.var(ddof=1)
    within = chain_var.mean()
    return float(np.sqrt((between / within + n - 1) / n))


def rhat_rank(ary):
    s = split_chains(ary)
    bulk = rhat_plain(z_scale(s))
    folded = np.abs(s - np.median(s))
    tail = rhat_plain(z_scale(folded))
    return max(bulk, tail)


def geyer_ess(mean_acov, var_chain_means, n_chain, n_draw):
    """ESS from the chain-averaged biased autocovariance mean_acov[t] and the variance (ddof=1) of the
    chain means — everything after the cross-chain reduction (ArviZ `_ess`)."""
    mean_var = mean_acov[0] * n_draw / (n_draw - 1.0)
    var_plus = mean_var * (n_draw - 1.0) / n_draw
    if n_chain > 1:
        var_plus += var_chain_means
    rho = np.zeros(n_draw)
    rho_even = 1.0
    rho[0] = rho_even
    rho_odd = 1.0 - (mean_var - mean_acov[1]) / var_plus
    rho[1] = rho_odd
    t = 1
    while t < (n_draw - 3) and (rho_even + rho_odd) > 0.0:
        rho_even = 1.0 - (mean_var - mean_acov[t + 1]) / var_plus
        rho_odd = 1.0 - (mean_var - mean_acov[t + 2]) / var_plus
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
    tau = -1.0 + 2.0 * np.sum(rho[: max_t + 1]) + np.sum(rho[max_t + 1: max_t + 2])
    tau = max(tau, 1.0 / np.log10(ess))
    return float("nan") if np.isnan(rho).any() else float(ess / tau)


def ess_plain(ary):
    ary = np.asarray(ary, dtype=float)
    if (ary.max() - ary.min()) < np.finfo(float).resolution:
        return float(ary.size)
    n_chain, n_draw = ary.shape
    acov = autocov(ary)
    vm = ary.mean(axis=1).var(ddof=1) if n_chain > 1 else 0.0
    return geyer_ess(acov.mean(axis=0), vm, n_chain, n_draw)


def ess_bulk(ary):
    return ess_plain(z_scale(split_chains(ary)))


def ess_quantile(ary, prob):
    q = np.quantile(np.asarray(ary), prob)
    return ess_plain(split_chains(np.asarray(ary) <= q))


def ess_tail(ary):
    return min(ess_quantile(ary, 0.05), ess_quantile(ary, 0.95))


def ess_mean(ary):
    return ess_plain(split_chains(ary))


def mcse_mean(ary):
    return float(np.std(ary, ddof=1) / np.sqrt(ess_mean(ary)))


def compute_diagnostics(samples):
    """samples (n_samples, n_chains, n_dim) -> the dict of run_benchmarks.py:161-195 (summary as arrays,
    unrounded)."""
    S, C, D = samples.shape
    x = np.asarray(samples).transpose(1, 0, 2)
    rh = np.array([rhat_rank(x[:, :, d]) for d in range(D)])
    eb = np.array([ess_bulk(x[:, :, d]) for d in range(D)])
    et = np.array([ess_tail(x[:, :, d]) for d in range(D)])
    mc = np.array([mcse_mean(x[:, :, d]) for d in range(D)])
    return {
        "rhat_max": float(rh.max()), "rhat_mean": float(rh.mean()),
        "ess_bulk_min": float(eb.min()), "ess_bulk_mean": float(eb.mean()),
        "ess_tail_min": float(et.min()), "ess_tail_mean": float(et.mean()),
        "mean_estimate": x.mean(axis=(0, 1)), "std_estimate": x.std(axis=(0, 1)),
        "rhat": rh, "ess_bulk": eb, "ess_tail": et,
        "summary": {"mean": x.mean(axis=(0, 1)), "sd": x.reshape(-1, D).std(axis=0, ddof=1), "mcse_mean": mc},
    }


def sliced_wasserstein_np(samples1, samples2, directions):
    """benchmarks/metrics.py:12-84 for given unit directions (n_proj, dim), equal sample counts."""
    p1 = np.sort(samples1 @ directions.T, axis=0)
    p2 = np.sort(samples2 @ directions.T, axis=0)
    if p1.shape[0] != p2.shape[0]:
        nq = min(p1.shape[0], p2.shape[0])
        qs = np.linspace(0, 1, nq)
        p1 = np.quantile(p1, qs, axis=0)
        p2 = np.quantile(p2, qs, axis=0)
    return float(np.mean(np.sqrt(np.mean((p1 - p2) ** 2, axis=0))))
