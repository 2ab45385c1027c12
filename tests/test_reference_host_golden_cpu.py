"""Host-side API and the oracle's target / schedule functions against outputs of the REFERENCE'S OWN SOURCE
(tests/golden/reference_host_golden.npz, made by tests/golden/make_reference_host_golden.py): dual averaging, joint
dual averaging, Welford, build_schedule, friction schedules, and every target's metadata, value and gradient (the
reference's gradient is jax.value_and_grad of the log-density; here it is the analytic one)."""
import ast
import os

import numpy as np
import pytest

from helpers import relerr
from mcmc_b200 import _abi
from mcmc_b200.benchmarks import targets as T
from mcmc_b200.tuning import adaptation as A, dual_averaging as DA, welford as W

BLOB = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "reference_host_golden.npz"))
TARGET_KEYS = [str(k) for k in BLOB["target_keys"]]


def parse_target_key(key):
    _, name, D, kw = key.split("|")
    return name, int(D), dict(ast.literal_eval(kw))


@pytest.mark.parametrize("tag", ["da_hmc", "da_nuts", "da_rwmh"])
def test_dual_averaging_matches_reference_source(tag):
    eps0, target = BLOB[f"{tag}|init"]
    want = BLOB[f"{tag}|states"]
    st = DA.da_init(float(eps0))
    rows = [[st.log_step, st.log_step_bar, st.H_bar, st.mu, st.count]]
    for i, a in enumerate(BLOB[f"{tag}|accept"]):
        if i in (12, 30):
            st = DA.da_reset(st)
            rows.append([st.log_step, st.log_step_bar, st.H_bar, st.mu, st.count])
        st = DA.da_update(st, float(a), float(target))
        rows.append([st.log_step, st.log_step_bar, st.H_bar, st.mu, st.count])
    got = np.array(rows, dtype=np.float64)
    assert got.shape == want.shape
    assert np.array_equal(got[:, 4], want[:, 4])
    assert relerr(got, want) <= 1e-13


def test_joint_dual_averaging_matches_reference_source():
    st = DA.joint_da_init(np.array([0.1, 1.0]))
    rows = []
    for a in BLOB["joint_da|accept"]:
        st = DA.joint_da_update(st, float(a), 0.65)
        rows.append(np.concatenate([np.asarray(st.log_params, dtype=np.float64), np.asarray(st.log_params_bar, dtype=np.float64),
                                    [st.H_bar, st.count]]))
    assert relerr(np.array(rows), BLOB["joint_da|states"]) <= 1e-13


def test_welford_matches_reference_source():
    import torch
    x = BLOB["welford|x"]
    st = W.welford_init(7)
    for row in x[:20]:
        st = W.welford_update(st, torch.from_numpy(row))
    assert relerr(np.stack([st.mean.numpy(), st.m2.numpy()]), BLOB["welford|after20"]) <= 1e-14
    st = W.welford_update_batch(st, torch.from_numpy(x[20:]))
    mean, var = W.welford_covariance(st)
    assert relerr(np.stack([st.mean.numpy(), st.m2.numpy(), mean.numpy(), var.numpy()]), BLOB["welford|final"]) <= 1e-14
    assert float(st.count) == float(BLOB["welford|count"].ravel()[0])
    _, v1 = W.welford_covariance(W.welford_update(W.welford_init(7), torch.from_numpy(x[0])))
    assert np.array_equal(v1.numpy(), BLOB["welford|one_sample_var"])


@pytest.mark.parametrize("tag,kw", [("default", {}), ("short", dict(exploration_steps=100, adaptation_windows=[25, 50], cooldown_steps=25))])
def test_build_schedule_matches_reference_source(tag, kw):
    s = A.build_schedule(**kw)
    got = np.array([[a, b, {"exploration": 0, "adaptation": 1, "cooldown": 2}[p]] for a, b, p in s])
    assert np.array_equal(got, BLOB[f"schedule|{tag}"])


FRICTION_KEYS = [k for k in BLOB.files if k.startswith("friction|")]


@pytest.mark.parametrize("key", FRICTION_KEYS)
def test_oracle_friction_schedule_matches_reference_source(oracle, key):
    _, name, L, eps, gam, steep = key.split("|")
    s, g = oracle.friction_table(_abi.SCHEDULE_IDS[name], int(L), float(eps), float(gam), float(steep))
    want = BLOB[key]
    assert relerr(g, want) <= 1e-14
    assert relerr(s, np.exp(-want * (0.5 * float(eps)))) <= 1e-14   # GRAHMC.py:193-194


@pytest.mark.parametrize("key", TARGET_KEYS)
def test_target_metadata_matches_reference_source(key):
    name, D, kw = parse_target_key(key)
    t = getattr(T, name)(dim=D, **kw)
    for field in ("true_mean", "true_cov"):
        want = BLOB[f"{key}|{field}"]
        got = getattr(t, field)
        if want.size == 0:
            assert got is None
        else:
            assert relerr(np.asarray(got, dtype=np.float64), want) <= 1e-12, field


@pytest.mark.parametrize("key", TARGET_KEYS)
def test_oracle_target_value_and_gradient_match_reference_autodiff(oracle, key):
    name, D, kw = parse_target_key(key)
    tag = T.resolve_log_prob(getattr(T, name)(dim=D, **kw).log_prob_fn)
    x = BLOB[key + "|x"]
    p = oracle.make_params(tag.target_id, tag.target_params, x.shape[0], D, 0)
    lp, g = oracle.target_eval(p, x, cov_inv=tag.cov_inv)
    assert relerr(lp, BLOB[key + "|lp"]) <= 1e-12
    assert relerr(g, BLOB[key + "|grad"]) <= 1e-11


def test_known_answers_from_reference_source():
    assert float(BLOB["kat|rosenbrock_at_ones"].ravel()[0]) == 0.0
    assert abs(float(BLOB["kat|standard_normal_at_zero"].ravel()[0]) + 10.0 * np.log(2.0 * np.pi)) < 1e-13


@pytest.mark.parametrize("tag,D", [("pass", 20), ("fail", 20), ("d10", 10)])
def test_z_test_gate_matches_reference_source(tag, D):
    """run_benchmarks.py:108-158 (Bonferroni threshold from scipy's norm.ppf there, restated here)."""
    from mcmc_b200 import diagnostics as DG
    tgt = T.ill_conditioned_gaussian(dim=D)
    r = DG.check_summary_statistics({"summary": {"mean": BLOB[f"ztest|{tag}|means"], "mcse_mean": BLOB[f"ztest|{tag}|mcse"]}}, tgt)
    want = BLOB[f"ztest|{tag}|out"]
    assert float(r["pass"]) == want[0]
    assert abs(r["max_z"] - want[1]) <= 1e-12 * max(1.0, want[1])
    assert abs(r["threshold"] - want[2]) <= 1e-9
    assert relerr(r["z_scores"], BLOB[f"ztest|{tag}|z"]) <= 1e-12


def test_gate_constants_match_reference_source():
    from mcmc_b200 import diagnostics as DG
    for k in BLOB.files:
        if k.startswith("gate_const|"):
            assert float(getattr(DG, k.split("|")[1])) == float(BLOB[k].ravel()[0]), k
