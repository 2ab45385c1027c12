"""Host-side tuning logic (tuning/dual_averaging.py, welford.py, adaptation.py of the reference):
known-answer facts from SURVEY.md §8c and the windowed warmup driven through the oracle stand-ins."""
import math

import numpy as np
import pytest
import torch

from mcmc_b200 import _abi
from mcmc_b200.benchmarks import targets as T
from mcmc_b200.tuning import (build_schedule, da_init, da_reset, da_update, joint_da_init, joint_da_update,
                              run_adaptive_warmup, welford_covariance, welford_init, welford_update,
                              welford_update_batch, dual_averaging_tune_hmc, dual_averaging_tune_rwmh)
from helpers import patch_tuning_to_oracle


def test_build_schedule_totals_and_da_updates():
    sch = build_schedule(2500)
    assert sch[0] == (0, 500, "exploration") and sch[-1] == (2375, 2500, "cooldown")
    assert [e - s for s, e, t in sch if t == "adaptation"] == [25, 50, 100, 200, 500, 1000]
    assert sch[-1][1] == 2500
    # adaptation.py:164-165 with update_freq=100 -> 26 dual-averaging updates over the 2500 steps
    assert sum(max(1, (e - s) // 100) for s, e, _ in sch) == 26
    assert build_schedule(None, 10, [5, 7], 3) == [(0, 10, "exploration"), (10, 15, "adaptation"),
                                                   (15, 22, "adaptation"), (22, 25, "cooldown")]


def test_dual_averaging_known_answers():
    s0 = da_init(0.1)
    assert s0.mu == pytest.approx(math.log(0.1)) and s0.log_step_bar == s0.log_step and s0.count == 0
    s1 = da_update(s0, 0.5, 0.65)
    # eta_1 = 1/11, H = 0.15/11, log eps = mu - (1/0.05) H; the first update sets log_bar = log_step
    assert s1.H_bar == pytest.approx(0.15 / 11)
    assert s1.log_step == pytest.approx(math.log(0.1) - 20.0 * 0.15 / 11)
    assert s1.log_step_bar == s1.log_step
    s2 = da_update(s1, 0.9, 0.65)
    H2 = (1 - 1 / 12) * s1.H_bar + (1 / 12) * (0.65 - 0.9)
    ls2 = s0.mu - (math.sqrt(2) / 0.05) * H2
    assert s2.log_step == pytest.approx(ls2)
    assert s2.log_step_bar == pytest.approx(2 ** -0.75 * ls2 + (1 - 2 ** -0.75) * s1.log_step_bar)
    r = da_reset(s2)
    assert r.mu == s2.log_step_bar and r.log_step == s2.log_step_bar and r.count == 0 and r.H_bar == 0.0
    assert da_reset(s0).mu == s0.log_step
    # accept above target -> step grows
    assert da_update(s0, 0.95, 0.65).log_step > s0.log_step


def test_joint_dual_averaging_clips_gamma():
    s = joint_da_init([0.1, 1.0])
    for _ in range(50):
        s = joint_da_update(s, 1.0, 0.65)   # always accepting -> both grow, gamma capped at 20
    assert s.log_params[1] == pytest.approx(math.log(20.0))
    assert s.log_params[0] > math.log(20.0) - 5 and s.log_params[0] > math.log(0.1)


def test_welford_matches_numpy():
    x = np.random.default_rng(0).standard_normal((37, 5)) * [1, 2, 3, 4, 5] + 3.0
    st = welford_init(5)
    for row in x[:10]:
        st = welford_update(st, row)
    st = welford_update_batch(st, x[10:])
    mean, var = welford_covariance(st)
    np.testing.assert_allclose(mean.numpy(), x.mean(0), rtol=1e-13)
    np.testing.assert_allclose(var.numpy(), x.var(0, ddof=1), rtol=1e-12)
    assert float(st.count) == 37.0
    # count <= 1: denominator clamps at 1 (welford.py:97-98)
    m, v = welford_covariance(welford_update(welford_init(2), [1.0, 2.0]))
    assert v.tolist() == [0.0, 0.0]


def test_warmup_learns_diagonal_metric(monkeypatch, oracle):
    """run_adaptive_warmup on an ill-conditioned Gaussian: the learned inv_mass_matrix approaches the
    true variances (shrunk to 1 with weight 5/(n+5)) and the step size ends near the target acceptance."""
    patch_tuning_to_oracle(monkeypatch)
    D, C = 6, 64
    tgt = T.ill_conditioned_gaussian(D, condition_number=25.0)
    q0 = np.random.default_rng(1).standard_normal((C, D))
    eps, w, pos, info = run_adaptive_warmup("hmc", tgt.log_prob_fn, None, q0, key=7, num_warmup=2500, num_steps=8,
                                            verbose=False)
    true_var = np.linspace(1.0, 25.0, D)
    n = 1000.0
    expect = n / (n + 5) * true_var + 5 / (n + 5)
    assert w.shape == (D,) and pos.shape == (C, D)
    np.testing.assert_allclose(w.numpy(), expect, rtol=0.15)
    assert info["num_da_updates"] == 26 and info["mass_matrix_learned"]
    # with the learned metric the target is ~isotropic: eps of order 1 for D=6, L=8
    assert 0.3 < eps < 2.5
    # the tuned step size reaches the target acceptance
    from helpers import oracle_sample_batch
    out = oracle_sample_batch(_abi.SAMPLER_HMC, tgt.log_prob_fn, pos, 99, 200, step_size=eps, num_steps=8,
                              inv_mass_matrix=w.numpy())
    acc = float(out["accept_sum"]) / out["accept_terms"]
    assert abs(acc - 0.65) < 0.12, acc


def test_warmup_identity_metric_and_errors(monkeypatch, oracle):
    patch_tuning_to_oracle(monkeypatch)
    tgt = T.standard_normal(4)
    q0 = np.zeros((8, 4))
    eps, w, pos, info = run_adaptive_warmup("grahmc", tgt.log_prob_fn, None, q0, key=1, num_warmup=2500,
                                            schedule_type="tanh", learn_mass_matrix=False, num_steps=4,
                                            tune_friction=False, verbose=False)
    assert torch.equal(w, torch.ones(4, dtype=torch.float64))
    assert info["gamma"] == 1.0 and info["steepness"] == 0.5 and not info["mass_matrix_learned"]
    with pytest.raises(ValueError):
        run_adaptive_warmup("gibbs", tgt.log_prob_fn, None, q0, key=1)
    with pytest.raises(KeyError):
        run_adaptive_warmup("grahmc", tgt.log_prob_fn, None, q0, key=1, schedule_type="cubic", verbose=False)


def test_tuners_converge(monkeypatch, oracle):
    patch_tuning_to_oracle(monkeypatch)
    tgt = T.standard_normal(5)
    q0 = np.random.default_rng(2).standard_normal((32, 5))
    eps, h = dual_averaging_tune_hmc(3, tgt.log_prob_fn, q0, num_steps=5, max_iter=150, min_iter=20)
    assert abs(np.mean(h["accept_history"][-20:]) - 0.65) < 0.1 and 0.2 < eps < 2.0
    sc, h = dual_averaging_tune_rwmh(3, tgt.log_prob_fn, q0, max_iter=150, min_iter=20)
    assert abs(np.mean(h["accept_history"][-20:]) - 0.234) < 0.08
    assert 0.5 < sc < 2.0   # 2.38/sqrt(5) = 1.06 is the asymptotic optimum
