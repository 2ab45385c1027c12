"""Pins the NumPy restatement of the ArviZ diagnostics (oracle/diagnostics_np.py) with known-answer cases,
and checks the host-side finish of mcmc_b200/diagnostics.py (Geyer truncation, R-hat, gates) against it."""
import numpy as np
import pytest

from mcmc_b200 import diagnostics as DG
from mcmc_b200.benchmarks import targets as T
from oracle import diagnostics_np as O


def _ar1(rng, chains, n, rho):
    y = np.zeros((chains, n))
    y[:, 0] = rng.standard_normal(chains)
    e = rng.standard_normal((chains, n)) * np.sqrt(1 - rho ** 2)
    for t in range(1, n):
        y[:, t] = rho * y[:, t - 1] + e[:, t]
    return y


def test_oracle_iid_known_answers():
    rng = np.random.default_rng(0)
    x = rng.standard_normal((16, 1000))
    n = x.size
    assert abs(O.ess_bulk(x) / n - 1) < 0.08
    assert abs(O.ess_tail(x) / n - 1) < 0.12
    assert abs(O.ess_mean(x) / n - 1) < 0.08
    assert abs(O.rhat_rank(x) - 1.0) < 0.003
    assert O.mcse_mean(x) == pytest.approx(1 / np.sqrt(n), rel=0.06)


@pytest.mark.parametrize("rho", [0.5, 0.9])
def test_oracle_ar1_ess(rho):
    y = _ar1(np.random.default_rng(1), 16, 8000, rho)
    theory = y.size * (1 - rho) / (1 + rho)
    assert abs(O.ess_mean(y) / theory - 1) < 0.12


def test_oracle_rhat_detects_shift_and_scale():
    rng = np.random.default_rng(2)
    x = rng.standard_normal((8, 500))
    shifted = x.copy()
    shifted[0] += 2.0
    assert O.rhat_rank(shifted) > 1.1
    scaled = x.copy()
    scaled[0] *= 5.0                       # same location, different scale: only the folded (tail) R-hat sees it
    assert O.rhat_plain(O.z_scale(O.split_chains(scaled))) < 1.02 < O.rhat_rank(scaled)
    assert O.z_scale(np.array([[1.0, 2.0], [2.0, 3.0]]))[0, 1] == O.z_scale(np.array([[1.0, 2.0], [2.0, 3.0]]))[1, 0]


def test_host_finish_matches_oracle():
    rng = np.random.default_rng(3)
    for rho in (0.0, 0.6, 0.95):
        y = O.split_chains(_ar1(rng, 6, 400, rho))
        n_chain, n_draw = y.shape
        acov = O.autocov(y)
        means = y.mean(axis=1)
        stats = (float(n_chain), float(means.mean()), float(((means - means.mean()) ** 2).sum()))
        want = O.ess_plain(y)
        got = DG._ess_from(acov.sum(axis=0), stats, n_draw)
        assert got == pytest.approx(want, rel=1e-12)
        assert DG._rhat_from(acov[:, 0].sum(), stats, n_draw) == pytest.approx(O.rhat_plain(y), rel=1e-12)
    # constant input (e.g. an indicator that is all ones): ArviZ returns the number of draws
    assert DG._ess_from(np.zeros(50), (4.0, 1.0, 0.0), 50) == 200.0
    # truncated autocovariance that ends before Geyer's sequence does must fail loudly
    y = O.split_chains(_ar1(rng, 4, 400, 0.99))
    acov = O.autocov(y)
    means = y.mean(axis=1)
    with pytest.raises(RuntimeError):
        DG._ess_from(acov.sum(axis=0)[:8], (float(y.shape[0]), 0.0, float(((means - means.mean()) ** 2).sum())), y.shape[1])


def test_z_threshold_and_gates():
    tgt = T.standard_normal(20)
    good = {"summary": {"mean": np.full(20, 0.01), "mcse_mean": np.full(20, 0.01)}}
    r = DG.check_summary_statistics(good, tgt)
    assert r["threshold"] == pytest.approx(3.0233, abs=1e-4) and r["pass"] and r["max_z"] == pytest.approx(1.0, rel=1e-5)
    bad = {"summary": {"mean": np.r_[np.zeros(19), 0.05], "mcse_mean": np.full(20, 0.01)}}
    assert not DG.check_summary_statistics(bad, tgt)["pass"]
    # the reference's literal inputs are az.summary's 3-decimal display values: tiny mcse rounds to 0 and fails
    tiny = {"summary": {"mean": np.full(20, 1e-4), "mcse_mean": np.full(20, 2.5e-4)}}
    assert DG.check_summary_statistics(tiny, tgt)["pass"]
    assert DG.check_summary_statistics(tiny, tgt, arviz_rounding=True)["pass"]        # 0.000 / (0.000 + 1e-8)
    assert DG.check_summary_statistics(T.rosenbrock(20) and bad, T.rosenbrock(20))["reason"] == "No ground truth"
    d = dict(rhat_max=1.005, ess_bulk_min=900.0, ess_tail_min=500.0)
    g = DG.quality_gates(d, num_samples=1000)
    assert g["usable"] and g["quality_pass"] and g["is_high_efficiency"]
    assert not DG.quality_gates(dict(d, rhat_max=1.02), 1000)["quality_pass"]
    assert DG.quality_gates(dict(d, rhat_max=1.02), 1000)["usable"]
    assert not DG.quality_gates(dict(d, ess_bulk_min=300.0), 1000)["usable"]
    assert not DG.quality_gates(d, 1000, divergence_rate=0.06)["usable"]
    assert not DG.quality_gates(d, 1000, stats_pass=False)["quality_pass"]


def test_reference_samplers_moments():
    for name, kw in (("standard_normal", {}), ("correlated_gaussian", {"correlation": 0.9}),
                     ("ill_conditioned_gaussian", {"condition_number": 100.0}), ("neals_funnel", {}),
                     ("gaussian_mixture", {"separation": 5.0}), ("log_gamma", {})):
        tgt = T.get_target(name, 6)
        x = T.get_reference_sampler(name, 6, **kw)(5, 40000)
        assert x.shape == (40000, 6)
        np.testing.assert_allclose(x.mean(0), tgt.true_mean, atol=0.15 * np.sqrt(np.diag(tgt.true_cov)).max())
        if name != "neals_funnel":           # funnel variances (e^4.5) converge too slowly for a moment check
            np.testing.assert_allclose(x.var(0), np.diag(tgt.true_cov), rtol=0.08)
    assert T.get_reference_sampler("rosenbrock", 20) is None and T.has_reference_sampler("rosenbrock")
    assert T.get_reference_sampler("unknown", 3) is None


def test_sw2_oracle_identity_and_shift():
    rng = np.random.default_rng(4)
    a = rng.standard_normal((5000, 4))
    dirs = rng.standard_normal((50, 4))
    dirs /= np.linalg.norm(dirs, axis=1, keepdims=True)
    assert O.sliced_wasserstein_np(a, a, dirs) == 0.0
    # a pure shift by v has 1-D W2 = |<v, dir>| along every direction
    v = np.array([1.0, 0, 0, 0])
    assert O.sliced_wasserstein_np(a, a + v, dirs) == pytest.approx(np.mean(np.abs(dirs @ v)), rel=1e-9)
