"""Pins the CPU oracle (oracle/mcmc_oracle.c).

The reference ships no golden vectors and jax is not installable
here; the oracle is pinned against the reference's own source in
tests/test_reference_golden_cpu.py (reference files run on a torch-fp64 stand-in
for jax).  This file holds the other pins:
  * Philox4x32-10 against the published Random123 known-answer vectors,
  * the known-answer facts derived from the reference source,
  * every gradient against torch fp64 autograd of the log-density as coded,
  * the C oracle against the independent NumPy restatement (oracle/oracle_np.py),
  * the committed golden fixtures (tests/golden/, made by tests/golden/make_golden.py).
"""
import math
import os

import numpy as np
import pytest
import torch

from helpers import TARGET_CASES, relerr, tid
from mcmc_b200 import _abi
from oracle import oracle_np as N

SCHEDS = ["constant", "tanh", "sigmoid", "linear", "sine"]


def test_philox_known_answers(oracle):
    # Random123 kat_vectors, philox4x32 10 rounds
    kat = [
        ([0, 0, 0, 0], [0, 0], [0x6627e8d5, 0xe169c58d, 0xbc57ac4c, 0x9b00dbd8]),
        ([0xffffffff] * 4, [0xffffffff] * 2, [0x408f276d, 0x41c83b0e, 0xa20bc7c6, 0x6d5451fd]),
        ([0x243f6a88, 0x85a308d3, 0x13198a2e, 0x03707344], [0xa4093822, 0x299f31d0],
         [0xd16cfe09, 0x94fdcceb, 0x5001e420, 0x24126ea1]),
    ]
    from mcmc_b200 import random as R
    for ctr, key, want in kat:
        assert [int(x) for x in oracle.philox(ctr, key)] == want
        got = R.philox4x32_10(*[np.array([c], dtype=np.uint32) for c in ctr], key[0], key[1])
        assert [int(g[0]) for g in got] == want


def _torch_logp(name, x, params):
    """The reference's log_prob_fn bodies, transcribed op-for-op to torch (1-D branch)."""
    D = x.shape[-1]
    two_pi = 2.0 * math.pi
    if name == "standard_normal":
        return -0.5 * (torch.sum(x ** 2) + D * math.log(two_pi))
    if name == "correlated_gaussian":
        rho = params[0]
        a = 1.0 / (1.0 - rho)
        b = -rho / ((1.0 - rho) * (1.0 + (D - 1) * rho))
        ci = a * torch.eye(D, dtype=x.dtype) + b * torch.ones((D, D), dtype=x.dtype)
        ld = (D - 1) * math.log(1.0 - rho) + math.log(1.0 + (D - 1) * rho)
        return -0.5 * (x @ ci @ x + ld + D * math.log(two_pi))
    if name == "ill_conditioned_gaussian":
        ev = torch.linspace(1.0, params[0], D, dtype=x.dtype)
        ci = torch.diag(1.0 / ev)
        return -0.5 * (x @ ci @ x + torch.sum(torch.log(ev)) + D * math.log(two_pi))
    if name == "neals_funnel":
        x0, xr = x[0], x[1:]
        lp0 = -0.5 * (x0 ** 2 / 9.0 + math.log(two_pi * 9.0))
        var = torch.exp(x0)
        return lp0 + -0.5 * (torch.sum(xr ** 2) / var + (D - 1) * x0 + (D - 1) * math.log(two_pi))
    if name == "log_gamma":
        al, be = params[0], params[1]
        valid = torch.all(x > 0)
        ln = math.lgamma(al) + al * math.log(be)
        pdf = (al - 1.0) * torch.log(torch.clamp(x, min=1e-10)) - be * x - ln
        return torch.where(valid, torch.sum(pdf), torch.tensor(-math.inf, dtype=x.dtype))
    if name == "student_t":
        df = params[0]
        ln = math.lgamma((df + 1.0) / 2.0) - math.lgamma(df / 2.0) - 0.5 * math.log(df * math.pi)
        return torch.sum(ln + -((df + 1.0) / 2.0) * torch.log(1.0 + x ** 2 / df))
    if name == "rosenbrock":
        a = 1.0 / params[0] ** 2
        xc, xn = x[:-1], x[1:]
        return -torch.sum((1.0 - xc) ** 2 + a * (xn - xc ** 2) ** 2)
    if name == "gaussian_mixture":
        sep = params[0]
        x0, xr = x[0], x[1:]
        m1 = -0.5 * (x0 + sep / 2.0) ** 2
        m2 = -0.5 * (x0 - sep / 2.0) ** 2
        mx = torch.maximum(m1, m2)
        lp0 = math.log(0.5) + mx + torch.log(torch.exp(m1 - mx) + torch.exp(m2 - mx)) - 0.5 * math.log(two_pi)
        return lp0 + -0.5 * (torch.sum(xr ** 2) + (D - 1) * math.log(two_pi))
    raise KeyError(name)


@pytest.mark.parametrize("name", list(TARGET_CASES))
@pytest.mark.parametrize("D", [2, 10, 20])
def test_target_value_and_grad_vs_autograd(oracle, name, D):
    params, _, init = TARGET_CASES[name]
    rng = np.random.default_rng(11)
    q = init(rng, 16, D)
    p = oracle.make_params(tid(name), params, 16, D, 0)
    lp, g = oracle.target_eval(p, q)
    lp_np, g_np = N.TARGETS[name](q) if name != "standard_normal" else N.standard_normal(q, True)
    for i in range(q.shape[0]):
        x = torch.tensor(q[i], dtype=torch.float64, requires_grad=True)
        v = _torch_logp(name, x, params)
        (gr,) = torch.autograd.grad(v, x)
        assert relerr(lp[i], v.item()) < 1e-12
        assert relerr(g[i], gr.numpy()) < 1e-11, (name, i)
    assert relerr(lp, lp_np) < 1e-12 and relerr(g, g_np) < 1e-12


def test_log_gamma_invalid_region(oracle):
    # x <= 0 anywhere: log p = -inf and the gradient is zero (where() kills the cotangent, targets.py:251)
    params = TARGET_CASES["log_gamma"][0]
    q = np.array([[1.0, -0.5, 2.0], [1.0, 2.0, 3.0]])
    p = oracle.make_params(tid("log_gamma"), params, 2, 3, 0)
    lp, g = oracle.target_eval(p, q)
    assert lp[0] == -np.inf and np.all(g[0] == 0.0)
    assert np.isfinite(lp[1]) and np.allclose(g[1], 1.0 / q[1] - 1.0)


def test_known_answer_facts(oracle):
    D = 20
    one = np.ones((1, D))
    zero = np.zeros((1, D))
    # rosenbrock log p(1,...,1) = 0 exactly (targets.py:356)
    lp, g = oracle.target_eval(oracle.make_params(tid("rosenbrock"), [0.1], 1, D, 0), one)
    assert lp[0] == 0.0 and np.all(g[0] == 0.0)
    # standard_normal log p(0) = -(D/2) log 2 pi
    lp, _ = oracle.target_eval(oracle.make_params(tid("standard_normal"), [1.0], 1, D, 0), zero)
    assert abs(lp[0] + 0.5 * D * math.log(2 * math.pi)) < 1e-13
    lp, _ = oracle.target_eval(oracle.make_params(tid("standard_normal"), [0.0], 1, D, 0), one)
    assert lp[0] == -0.5 * D  # README.md:51 variant
    # correlated gaussian: closed-form precision times covariance is the identity (targets.py:79-85)
    rho = 0.9
    a = 1.0 / (1.0 - rho)
    b = -rho / ((1.0 - rho) * (1.0 + (D - 1) * rho))
    P = a * np.eye(D) + b * np.ones((D, D))
    cov = (1 - rho) * np.eye(D) + rho * np.ones((D, D))
    assert np.allclose(P @ cov, np.eye(D), atol=1e-12)


@pytest.mark.parametrize("L", [1, 2, 7, 16, 128])
def test_friction_schedules(oracle, L):
    eps, gam = 0.05, 1.3
    for name in SCHEDS:
        steep = 0.5 if name == "tanh" else 2.0
        s, g = oracle.friction_table(_abi.SCHEDULE_IDS[name], L, eps, gam, steep)
        s_np, g_np = N.friction_table(name, L, eps, gam, steep)
        assert relerr(s, s_np) < 1e-15 and relerr(g, g_np) < 1e-15
        # values at i = 0
        want0 = {"constant": -gam if L > 0 else gam, "tanh": -gam * math.tanh(steep),
                 "sigmoid": gam * (2.0 / (1.0 + math.exp(steep / 2)) - 1.0), "linear": -gam, "sine": -gam}[name]
        assert abs(g[0] - want0) < 1e-14
        if name == "constant":  # sign flips at 2i >= L (GRAHMC.py:74)
            assert all((g[i] < 0) == (2 * i < L) for i in range(L))
        assert np.allclose(s, np.exp(-g * 0.5 * eps), rtol=1e-15)


@pytest.mark.parametrize("name", list(TARGET_CASES))
def test_c_oracle_matches_numpy_oracle_hmc_grahmc(oracle, name):
    params, eps, init = TARGET_CASES[name]
    rng = np.random.default_rng(3)
    C_, D, S, B, L = 5, 20, 4, 2, 8
    q0 = init(rng, C_, D)
    z = rng.standard_normal((B + S, C_, D))
    u = rng.random((B + S, C_))
    w = np.linspace(0.5, 2.0, D)
    tgt = N.TARGETS[name]
    for sched in [None] + SCHEDS:
        p = oracle.make_params(tid(name), params, C_, D, S, B, L, eps, gamma=0.7, steepness=0.5,
                               schedule_id=_abi.SCHEDULE_IDS.get(sched, 0), flags=_abi.FLAG_TRACK_PROPOSALS)
        r = oracle.run(_abi.SAMPLER_GRAHMC if sched else _abi.SAMPLER_HMC, p, q0, inv_mass=w, z=z, u=u)
        rn = N.hmc_run(tgt, q0, eps, L, S, z, u, burn_in=B, inv_mass=w, schedule=sched, gamma=0.7, steepness=0.5)
        for k in ("samples", "log_probs", "pre_q", "pre_lp", "prop_q", "prop_lp", "delta_h", "final_q", "final_lp", "final_g"):
            assert relerr(r[k], rn[k]) < 1e-11, (name, sched, k)
        assert np.array_equal(r["accept_count"], rn["accept_count"])
        assert np.array_equal(r["accept_rate"], rn["accept_rate"])
        assert int(r["counters"][0]) == int(rn["accept_count"].sum())


@pytest.mark.parametrize("name", ["standard_normal", "rosenbrock", "log_gamma", "neals_funnel"])
def test_c_oracle_matches_numpy_oracle_rwmh(oracle, name):
    params, eps, init = TARGET_CASES[name]
    rng = np.random.default_rng(4)
    C_, D, S, B = 7, 10, 12, 3
    q0 = init(rng, C_, D)
    z = rng.standard_normal((B + S, C_, D))
    u = rng.random((B + S, C_))
    scale = 0.3 if name != "rosenbrock" else 0.02
    p = oracle.make_params(tid(name), params, C_, D, S, B, 0, scale)
    r = oracle.run(_abi.SAMPLER_RWMH, p, q0, z=z, u=u)
    rn = N.rwmh_run(N.TARGETS[name], q0, scale, S, z, u, burn_in=B)
    for k in ("samples", "log_probs", "final_q", "final_lp"):
        assert relerr(r[k], rn[k]) < 1e-12, (name, k)
    assert np.array_equal(r["accept_count"], rn["accept_count"])


@pytest.mark.parametrize("name", ["standard_normal", "neals_funnel", "rosenbrock"])
def test_c_oracle_matches_numpy_oracle_nuts(oracle, name):
    params, eps, init = TARGET_CASES[name]
    rng = np.random.default_rng(5)
    C_, D, S, B, md = 3, 10, 4, 1, 6
    q0 = init(rng, C_, D)
    z = rng.standard_normal((B + S, C_, D))
    u = rng.random((B + S, C_))
    du = rng.random((B + S, C_, md))
    pu = rng.random((B + S, C_, md))
    w = np.linspace(0.8, 1.5, D)
    p = oracle.make_params(tid(name), params, C_, D, S, B, 0, eps, max_tree_depth=md)
    r = oracle.run(_abi.SAMPLER_NUTS, p, q0, inv_mass=w, z=z, u=u, dir_u=du, prop_u=pu)
    rn = N.nuts_run(N.TARGETS[name], q0, eps, S, z, u, du, pu, burn_in=B, inv_mass=w, max_tree_depth=md)
    for k in ("samples", "log_probs", "mean_accept_probs", "final_q", "final_lp", "final_g"):
        assert relerr(r[k], rn[k]) < 1e-10, (name, k)
    assert np.array_equal(r["tree_depths"], rn["tree_depths"])
    assert np.all(r["accept_count"] == S) and np.all(r["accept_rate"] == 1.0)  # NUTS.py:385


def test_leapfrog_reversibility_and_gamma_zero(oracle):
    # integrate, negate p, integrate -> back at the start; GRAHMC with gamma=0 == HMC
    rng = np.random.default_rng(6)
    C_, D, L, eps = 4, 20, 16, 0.1
    q0 = rng.standard_normal((C_, D))
    z = rng.standard_normal((1, C_, D))
    u = np.zeros((1, C_))  # log(0) = -inf: always accept
    p = oracle.make_params(tid("neals_funnel"), [], C_, D, 1, 0, L, eps, flags=_abi.FLAG_TRACK_PROPOSALS)
    fwd = oracle.run(_abi.SAMPLER_HMC, p, q0, z=z, u=u)
    # final momentum is not exported; recover it with the NumPy oracle's internals via a second forward run
    pg = oracle.make_params(tid("neals_funnel"), [], C_, D, 1, 0, L, eps, gamma=0.0, schedule_id=1,
                            flags=_abi.FLAG_TRACK_PROPOSALS)
    gr = oracle.run(_abi.SAMPLER_GRAHMC, pg, q0, z=z, u=u)
    assert np.array_equal(fwd["prop_q"], gr["prop_q"]) and np.array_equal(fwd["delta_h"], gr["delta_h"])
    # reversibility through the NumPy restatement (it exposes p)
    tgt = N.neals_funnel
    q, pp = q0.copy(), z[0].copy()
    lp, g = tgt(q)
    for sign in (1, 1):
        for _ in range(L):
            pp = pp + 0.5 * eps * g
            q = q + eps * pp
            lp, g = tgt(q)
            pp = pp + 0.5 * eps * g
        if sign == 1:
            pp = -pp
    assert relerr(q, q0) < 1e-11


def test_welford_fold_matches_reference_formula(oracle):
    rng = np.random.default_rng(8)
    C_, D, S = 3, 5, 9
    q0 = rng.standard_normal((C_, D))
    z = rng.standard_normal((S, C_, D))
    u = rng.random((S, C_))
    p = oracle.make_params(tid("standard_normal"), [1.0], C_, D, S, 0, 4, 0.3, flags=_abi.FLAG_ACCUM_WELFORD)
    r = oracle.run(_abi.SAMPLER_HMC, p, q0, z=z, u=u, welford=(np.zeros((C_, D)), np.zeros((C_, D))))
    smp = r["samples"]
    assert relerr(r["welford_mean"], smp.mean(axis=0)) < 1e-13
    assert relerr(r["welford_m2"], ((smp - smp.mean(axis=0)) ** 2).sum(axis=0)) < 1e-12
    sv = oracle.welford_finalize(r["welford_m2"], S)
    assert relerr(sv, smp.var(axis=0, ddof=1).sum(axis=0)) < 1e-12


def test_golden_fixtures(oracle):
    path = os.path.join(os.path.dirname(__file__), "golden", "hot_path_golden.npz")
    assert os.path.exists(path), "run tests/golden/make_golden.py"
    gold = np.load(path, allow_pickle=False)
    from golden.make_golden import CASES, run_case
    for case in CASES:
        got = run_case(oracle, case)
        for k, v in got.items():
            want = gold[f"{case['id']}/{k}"]
            assert relerr(v, want) < 1e-12, (case["id"], k)
