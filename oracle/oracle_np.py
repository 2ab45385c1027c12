"""NumPy restatement of the reference's samplers (second, independent oracle).

TEST INFRASTRUCTURE — imported only by tests/.  It exists to cross-check the C
oracle (oracle/mcmc_oracle.c): both are written from the reference source, in
different languages and loop orders (this one is vectorised over chains like
the reference's vmap; the C one is chain-outer), and must agree to rounding.

PARITY (see the header of mcmc_oracle.c): pinned against the reference's own source run on a
torch-fp64 stand-in for jax (tests/golden/reference_golden.npz, tests/test_reference_golden_cpu.py);
jax/XLA numerics and threefry themselves stay unpinned (jax is not installable here).

Follows, with injected randomness (z: (iters, C, D) normals, u: (iters, C)):
  samplers/HMC.py:86-134, 138-219, 259-314
  samplers/GRAHMC.py:69-106, 152-211, 215-261, 265-349, 398-456
  samplers/RWMH.py:73-119, 148-178
  samplers/NUTS.py:90-198, 202-396 (per chain, Python loops: small cases only)
  benchmarks/targets.py:50-52, 90-98, 130-137, 166-185, 236-251, 300-311, 367-387, 432-454
"""
from __future__ import annotations

import math
import numpy as np

LOG_2PI = math.log(2.0 * math.pi)


# --------------------------------------------------------------------- targets
# Each returns (log_prob (C,), grad (C, D)) for x of shape (C, D); the value is
# written exactly as the reference's batched branch, the gradient analytically.

def standard_normal(x, with_norm=True):
    D = x.shape[-1]
    s = np.sum(x ** 2, axis=-1)
    lp = -0.5 * (s + D * LOG_2PI) if with_norm else -0.5 * s
    return lp, -x


def correlated_gaussian(x, rho=0.9):
    D = x.shape[-1]
    a = 1.0 / (1.0 - rho)
    b = -rho / ((1.0 - rho) * (1.0 + (D - 1) * rho))
    cov_inv = a * np.eye(D) + b * np.ones((D, D))
    logdet = (D - 1) * np.log(1.0 - rho) + np.log(1.0 + (D - 1) * rho)
    y = x @ cov_inv
    maha = np.sum(y * x, axis=-1)
    return -0.5 * (maha + logdet + D * LOG_2PI), -y


def ill_conditioned_gaussian(x, kappa=100.0):
    D = x.shape[-1]
    lam = np.linspace(1.0, kappa, D)
    inv = 1.0 / lam
    maha = np.sum((x * inv) * x, axis=-1)
    return -0.5 * (maha + np.sum(np.log(lam)) + D * LOG_2PI), -(x * inv)


def neals_funnel(x):
    D = x.shape[-1]
    x0 = x[:, 0]
    xr = x[:, 1:]
    Dr = D - 1
    lp0 = -0.5 * (x0 ** 2 / 9.0 + np.log(2.0 * np.pi * 9.0))
    var = np.exp(x0)
    s = np.sum(xr ** 2, axis=-1)
    lpr = -0.5 * (s / var + Dr * x0 + Dr * LOG_2PI)
    g = np.empty_like(x)
    g[:, 0] = -x0 / 9.0 + 0.5 * s / var - 0.5 * Dr
    g[:, 1:] = -xr / var[:, None]
    return lp0 + lpr, g


def log_gamma(x, alpha=2.0, beta=1.0):
    lognorm = math.lgamma(alpha) + alpha * math.log(beta)
    valid = np.all(x > 0, axis=-1)
    xm = np.maximum(x, 1e-10)
    with np.errstate(invalid="ignore", divide="ignore"):
        pdf = (alpha - 1.0) * np.log(xm) - beta * x - lognorm
    res = np.sum(pdf, axis=-1)
    lp = np.where(valid, res, -np.inf)
    dmax = np.where(x > 1e-10, 1.0, np.where(x == 1e-10, 0.5, 0.0))
    g = dmax * (alpha - 1.0) / xm - beta
    g = np.where(valid[:, None], g, 0.0)
    return lp, g


def student_t(x, df=3.0):
    lognorm = math.lgamma((df + 1.0) / 2.0) - math.lgamma(df / 2.0) - 0.5 * math.log(df * math.pi)
    lk = -((df + 1.0) / 2.0) * np.log(1.0 + x ** 2 / df)
    return np.sum(lognorm + lk, axis=-1), -(df + 1.0) * x / (df + x ** 2)


def rosenbrock(x, scale=0.1):
    a = 1.0 / (scale ** 2)
    xc = x[:, :-1]
    xn = x[:, 1:]
    t = xn - xc ** 2
    U = np.sum((1.0 - xc) ** 2 + a * t ** 2, axis=-1)
    g = np.zeros_like(x)
    g[:, :-1] += 2.0 * (1.0 - xc) + 4.0 * a * xc * t
    g[:, 1:] += -2.0 * a * t
    return -U, g


def gaussian_mixture(x, separation=5.0):
    D = x.shape[-1]
    x0 = x[:, 0]
    xr = x[:, 1:]
    m1 = -0.5 * (x0 + separation / 2.0) ** 2
    m2 = -0.5 * (x0 - separation / 2.0) ** 2
    mx = np.maximum(m1, m2)
    e1, e2 = np.exp(m1 - mx), np.exp(m2 - mx)
    lp0 = np.log(0.5) + mx + np.log(e1 + e2) - 0.5 * LOG_2PI
    lpr = -0.5 * (np.sum(xr ** 2, axis=-1) + (D - 1) * LOG_2PI)
    g = np.empty_like(x)
    g[:, 0] = (e1 * (-(x0 + separation / 2.0)) + e2 * (-(x0 - separation / 2.0))) / (e1 + e2)
    g[:, 1:] = -xr
    return lp0 + lpr, g


def dense_gaussian(x, cov_inv, logdet):
    D = x.shape[-1]
    y = x @ cov_inv
    maha = np.sum(y * x, axis=-1)
    return -0.5 * (maha + logdet + D * LOG_2PI), -0.5 * (y + x @ cov_inv.T)


TARGETS = {
    "standard_normal": standard_normal,
    "correlated_gaussian": correlated_gaussian,
    "ill_conditioned_gaussian": ill_conditioned_gaussian,
    "neals_funnel": neals_funnel,
    "log_gamma": log_gamma,
    "student_t": student_t,
    "rosenbrock": rosenbrock,
    "gaussian_mixture": gaussian_mixture,
}


# ------------------------------------------------------------------- schedules

def friction_gamma(name, t, T, gamma, steepness):
    if name == "constant":
        return -gamma if t < T / 2 else gamma
    if name == "tanh":
        return gamma * math.tanh(steepness * (2.0 * t / T - 1.0))
    if name == "sigmoid":
        return gamma * (2.0 / (1.0 + math.exp(-(steepness * (t / T - 0.5)))) - 1.0)
    if name == "linear":
        return -gamma + (2.0 * gamma * t / T)
    if name == "sine":
        return gamma * math.sin(math.pi * (t / T - 0.5))
    raise KeyError(name)


def friction_table(name, L, eps, gamma, steepness):
    T = eps * L
    gam = np.array([friction_gamma(name, i * eps, T, gamma, steepness) for i in range(L)])
    return np.exp(-gam * (0.5 * eps)), gam


# -------------------------------------------------------------------- samplers

def _min_nan(a, b):
    return np.minimum(a, b)  # numpy minimum propagates NaN like jnp.minimum


def hmc_run(target, init_q, step_size, num_steps, num_samples, z, u, burn_in=0, inv_mass=None,
            schedule=None, gamma=0.0, steepness=1.0):
    """hmc_run (schedule None) or rahmc_run (schedule = name).  Returns a dict."""
    q = np.array(init_q, dtype=np.float64, copy=True)
    C, D = q.shape
    w = np.ones(D) if inv_mass is None else np.asarray(inv_mass, dtype=np.float64)
    lp, g = target(q)
    eps = float(step_size)
    half_eps = 0.5 * eps
    s_tab = friction_table(schedule, num_steps, eps, gamma, steepness)[0] if schedule else None
    acc = np.zeros(C, dtype=np.int32)
    out = {k: [] for k in ("samples", "log_probs", "pre_q", "pre_lp", "prop_q", "prop_lp", "delta_h", "accept")}
    for it in range(burn_in + num_samples):
        if it == burn_in:
            acc[:] = 0
        p = z[it] / np.sqrt(w)
        H0 = -lp + 0.5 * np.sum(p ** 2 * w, axis=-1)
        qn, gn, lpn = q.copy(), g.copy(), lp.copy()
        for i in range(num_steps):
            if s_tab is not None:
                p = p * s_tab[i]
            p = p + half_eps * gn
            qn = qn + eps * (p * w)
            lpn, gn = target(qn)
            p = p + half_eps * gn
            if s_tab is not None:
                p = p * s_tab[i]
        p = -p
        H1 = -lpn + 0.5 * np.sum(p ** 2 * w, axis=-1)
        H1 = np.where(np.isfinite(H1), H1, 1e10)
        with np.errstate(divide="ignore"):
            accept = np.log(u[it]) < _min_nan(0.0, H0 - H1)
        if it >= burn_in:
            out["pre_q"].append(q.copy()); out["pre_lp"].append(lp.copy())
            out["prop_q"].append(qn.copy()); out["prop_lp"].append(lpn.copy())
            out["delta_h"].append(H1 - H0); out["accept"].append(accept.copy())
        q = np.where(accept[:, None], qn, q)
        g = np.where(accept[:, None], gn, g)
        lp = np.where(accept, lpn, lp)
        acc += accept.astype(np.int32)
        if it >= burn_in:
            out["samples"].append(q.copy()); out["log_probs"].append(lp.copy())
    res = {k: np.array(v) for k, v in out.items()}
    if num_samples == 0:
        acc[:] = 0   # reset after burn-in even when nothing is sampled (HMC.py:276-282)
    res.update(accept_count=acc, accept_rate=acc.astype(np.float32) / np.float32(num_samples),
               final_q=q, final_lp=lp, final_g=g)
    return res


def rwmh_run(target, init_q, scale, num_samples, z, u, burn_in=0):
    q = np.array(init_q, dtype=np.float64, copy=True)
    C, D = q.shape
    lp = target(q)[0]
    acc = np.zeros(C, dtype=np.int32)
    samples, lps = [], []
    for it in range(burn_in + num_samples):
        if it == burn_in:
            acc[:] = 0
        prop = q + scale * z[it]
        lpn = target(prop)[0]
        with np.errstate(divide="ignore", invalid="ignore"):
            accept = np.log(u[it]) < _min_nan(0.0, lpn - lp)
        q = np.where(accept[:, None], prop, q)
        lp = np.where(accept, lpn, lp)
        acc += accept.astype(np.int32)
        if it >= burn_in:
            samples.append(q.copy()); lps.append(lp.copy())
    if num_samples == 0:
        acc[:] = 0
    return dict(samples=np.array(samples), log_probs=np.array(lps), accept_count=acc,
                accept_rate=acc.astype(np.float32) / np.float32(num_samples), final_q=q, final_lp=lp)


def nuts_run(target, init_q, step_size, num_samples, z, u, dir_u, prop_u, burn_in=0, inv_mass=None,
             max_tree_depth=10, delta_max=1000.0):
    """nuts_run; dir_u/prop_u: (iters, C, max_tree_depth) uniforms (direction = +1 iff dir_u < 0.5)."""
    q_all = np.array(init_q, dtype=np.float64, copy=True)
    C, D = q_all.shape
    w = np.ones(D) if inv_mass is None else np.asarray(inv_mass, dtype=np.float64)
    f1 = lambda x: tuple(a[0] for a in target(x[None, :]))
    lp_all, g_all = target(q_all)
    S = num_samples
    samples = np.zeros((S, C, D)); lps = np.zeros((S, C))
    depths = np.zeros((S, C), dtype=np.int32); maps = np.zeros((S, C))
    n_grad = 0
    energy = lambda lp, p: -lp + 0.5 * np.sum(p ** 2 * w)
    for c in range(C):
        q, lp, g = q_all[c].copy(), lp_all[c], g_all[c].copy()
        for it in range(burn_in + S):
            p0 = z[it, c] / np.sqrt(w)
            h0 = energy(lp, p0)
            with np.errstate(divide="ignore"):
                log_u = np.log(u[it, c]) - h0
            left = [q.copy(), p0.copy(), g.copy()]
            right = [q.copy(), p0.copy(), g.copy()]
            prop = [q.copy(), lp, g.copy()]
            n_valid, n_steps, sum_alpha, depth, diverged = 1, 0, 0.0, 0, False
            while True:
                dq = right[0] - left[0]
                u_turn = (np.dot(dq, left[1]) < 0) | (np.dot(dq, right[1]) < 0)
                if not ((depth < max_tree_depth) and (not u_turn) and (not diverged)):
                    break
                direction = 1 if dir_u[it, c, depth] < 0.5 else -1
                qs, ps, gs = [a.copy() for a in (left if direction == -1 else right)]
                num = 2 ** depth
                se = direction * step_size
                lps_ = np.nan
                sa = 0.0
                for _ in range(num):
                    ps = ps + 0.5 * se * gs
                    qs = qs + se * (ps * w)
                    lps_, gs = f1(qs)
                    ps = ps + 0.5 * se * gs
                    with np.errstate(over="ignore", invalid="ignore"):
                        sa = sa + np.exp(_min_nan(0.0, h0 - energy(lps_, ps)))
                n_grad += num
                h_new = energy(lps_, ps)
                in_slice = log_u <= -h_new
                is_div = (h_new - h0) > delta_max
                is_valid = bool(in_slice and not is_div)
                if direction == -1:
                    left = [qs, ps, gs]
                else:
                    right = [qs, ps, gs]
                n_new = num if is_valid else 0
                total = n_valid + n_new
                ap = np.float32(n_new / max(total, 1)) if (total > 0 and is_valid) else np.float32(0.0)
                if prop_u[it, c, depth] < float(ap):
                    prop = [qs.copy(), lps_, gs.copy()]
                n_valid = total
                sum_alpha += sa
                n_steps += num
                depth += 1
                diverged = diverged or bool(is_div)
            m = sum_alpha / max(n_steps, 1)
            if np.isnan(m) or np.isinf(m):
                m = 0.65
            q, lp, g = prop[0], prop[1], prop[2]
            if it >= burn_in:
                s = it - burn_in
                samples[s, c] = q; lps[s, c] = lp; depths[s, c] = depth; maps[s, c] = m
        q_all[c], lp_all[c], g_all[c] = q, lp, g
    return dict(samples=samples, log_probs=lps, tree_depths=depths, mean_accept_probs=maps,
                final_q=q_all, final_lp=lp_all, final_g=g_all, n_grad=n_grad)
