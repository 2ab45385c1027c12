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


@pytest.mark.parametrize("tag,sampler,kw", [("hmc", "hmc", {}), ("nuts", "nuts", {}), ("hmc_identity", "hmc", dict(learn_mass_matrix=False))])
def test_warmup_control_flow_matches_reference_source(monkeypatch, oracle, tag, sampler, kw):
    """run_adaptive_warmup with the sampler replaced by the same deterministic function on both sides
    (tests/golden/warmup_mock.py): 26 dual-averaging updates at the same step sizes, the same learned metric.
    The reference averages float32 accept rates (HMC.py:313, adaptation.py:218); here the integer counts are summed
    exactly, so step sizes agree to float32 rounding (1e-6) and the metric, which does not depend on them, to 1e-12."""
    import torch
    from golden.warmup_mock import mock_batch
    from helpers import oracle_welford_sum_var
    calls, steps = [0], []

    def fake_sample_batch(sid, log_prob_fn, position, key, num_samples, *, step_size, inv_mass_matrix=None, welford=None,
                          welford_count=0.0, **_k):
        pos = position.numpy()
        smp, cnt = mock_batch(calls[0], pos, step_size, None if inv_mass_matrix is None else inv_mass_matrix.numpy(), num_samples)
        calls[0] += 1
        steps.append(float(step_size))
        if welford is not None:   # what the kernels do under MCMC_FLAG_ACCUM_WELFORD: one fold per stored draw and chain
            mean, m2 = welford[0].numpy(), welford[1].numpy()
            n = float(welford_count)
            for x in smp:
                n += 1.0
                delta = x - mean
                mean += delta / n
                m2 += delta * (x - mean)
        out = {"final_q": torch.from_numpy(smp[-1].copy()), "counters": torch.tensor([int(cnt.sum()), 0, 0, 0])}
        if sid == _abi.SAMPLER_NUTS:   # adaptation.py:187: the statistic is the mean accept probability of the trajectories
            rate = (cnt.astype(np.float32) / np.float32(num_samples)).astype(np.float64)
            out["accept_sum"] = torch.tensor(float(rate.sum() * num_samples))
        else:
            out["accept_sum"] = torch.tensor(float(cnt.sum()))
        out["accept_terms"] = len(cnt) * int(num_samples)
        return out

    monkeypatch.setattr(A, "sample_batch", fake_sample_batch)
    monkeypatch.setattr(A, "welford_sum_var", oracle_welford_sum_var)
    monkeypatch.setattr(A, "_as_state_tensor", lambda x: torch.as_tensor(np.asarray(x), dtype=torch.float64).contiguous()
                        if not isinstance(x, torch.Tensor) else x.to(torch.float64).contiguous())
    eps, w, pos, info = A.run_adaptive_warmup(sampler, None, None, BLOB[f"warmup|{tag}|q0"], 1, num_warmup=2500, target_accept=0.65,
                                              verbose=False, **kw)
    want_steps = BLOB[f"warmup|{tag}|step_sizes"]
    assert len(steps) == len(want_steps) == 26 and info["num_da_updates"] == 26
    assert relerr(np.array(steps), want_steps) <= 2e-6
    assert abs(eps - float(BLOB[f"warmup|{tag}|final_step"].ravel()[0])) <= 2e-6 * eps
    assert relerr(w.numpy(), BLOB[f"warmup|{tag}|inv_mass"]) <= 1e-12
    assert relerr(pos.numpy(), BLOB[f"warmup|{tag}|position"]) <= 1e-12


@pytest.mark.parametrize("tag", ["rwmh", "hmc", "nuts"])
def test_tuner_convergence_loops_match_reference_source(monkeypatch, tag):
    """dual_averaging_tune_{rwmh,hmc,nuts} (dual_averaging.py:16-382) with the same mock sampler on both sides: same
    iteration of convergence, same histories (to the float32 rounding of the reference's accept rates)."""
    import torch
    from golden.warmup_mock import mock_batch
    calls = [0]

    def fake_sample_batch(sid, log_prob_fn, position, key, num_samples, *, step_size, **_k):
        pos = position.numpy() if isinstance(position, torch.Tensor) else np.asarray(position)
        smp, cnt = mock_batch(calls[0], pos, step_size * (0.25 if tag == "rwmh" else 1.0), None, num_samples)
        calls[0] += 1
        rate = (cnt.astype(np.float32) / np.float32(num_samples)).astype(np.float64)
        depth = 3 + calls[0] % 2
        return {"final_q": torch.from_numpy(smp[-1].copy()), "accept_terms": len(cnt) * int(num_samples),
                "accept_sum": torch.tensor(float(rate.sum() * num_samples) if tag == "nuts" else float(cnt.sum())),
                "counters": torch.tensor([int(cnt.sum()), 0, depth * len(cnt) * int(num_samples), 0])}

    monkeypatch.setattr(DA, "sample_batch", fake_sample_batch)
    q0 = torch.from_numpy(BLOB[f"tuner|{tag}|q0"])
    if tag == "rwmh":
        v, h = DA.dual_averaging_tune_rwmh(2, None, q0, max_iter=400)
        hist = h["scale_history"]
    elif tag == "hmc":
        v, h = DA.dual_averaging_tune_hmc(2, None, q0, 8, max_iter=400)
        hist = h["step_size_history"]
    else:
        v, h = DA.dual_averaging_tune_nuts(2, None, q0, max_iter=400)
        hist = h["step_size_history"]
        assert relerr(np.array(h["tree_depth_history"]), BLOB[f"tuner|{tag}|depth"]) <= 1e-12
    assert h["converged_iter"] == int(BLOB[f"tuner|{tag}|converged_iter"].ravel()[0])
    assert len(hist) == len(BLOB[f"tuner|{tag}|history"])
    assert relerr(np.array(hist), BLOB[f"tuner|{tag}|history"]) <= 5e-6
    assert relerr(np.array(h["accept_history"]), BLOB[f"tuner|{tag}|accept"]) <= 1e-6
    assert abs(v - float(BLOB[f"tuner|{tag}|value"].ravel()[0])) <= 5e-6 * v


def test_joint_tuner_matches_reference_source(monkeypatch):
    """joint_tune_grahmc (dual_averaging.py:668-804) with the same mock sampler on both sides."""
    import torch
    from golden.warmup_mock import mock_batch
    calls = [0]

    def fake_sample_batch(sid, log_prob_fn, position, key, num_samples, *, step_size, gamma, **_k):
        pos = position.numpy() if isinstance(position, torch.Tensor) else np.asarray(position)
        smp, cnt = mock_batch(calls[0], pos, step_size * (1.0 + 0.25 * gamma), None, num_samples)
        calls[0] += 1
        return {"final_q": torch.from_numpy(smp[-1].copy()), "accept_terms": len(cnt) * int(num_samples),
                "accept_sum": torch.tensor(float(cnt.sum())), "counters": torch.tensor([int(cnt.sum()), 0, 0, 0])}

    monkeypatch.setattr(DA, "sample_batch", fake_sample_batch)
    st, gm, sp, h = DA.joint_tune_grahmc(3, None, None, torch.from_numpy(BLOB["joint_tuner|q0"]), 8, schedule_type="tanh", max_iter=120)
    want = BLOB["joint_tuner|final"]
    assert abs(st - want[0]) <= 5e-6 * want[0] and abs(gm - want[1]) <= 5e-6 * want[1] and sp == want[2]
    assert relerr(np.array([h["step_size"], h["gamma"], h["accept_rate"]]), BLOB["joint_tuner|history"]) <= 5e-6
