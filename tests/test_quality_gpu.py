"""Statistical validation under the native Philox stream: the reference's own quality gates
(run_benchmarks.py:777-804: rank-normalised R-hat, bulk/tail ESS, Bonferroni z-test of the means, Sliced-W2)
for every sampler, plus the windowed warmup on the device."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

C, D = 2048, 20


def _gates(samples, tgt, name, num_samples, **kw):
    from mcmc_b200 import diagnostics as DG
    from mcmc_b200.benchmarks.metrics import compute_sliced_w2
    from mcmc_b200.benchmarks.targets import get_reference_sampler
    d = DG.compute_diagnostics(samples)
    z = DG.check_summary_statistics(d, tgt)
    g = DG.quality_gates(d, num_samples, stats_pass=z["pass"])
    sw2 = compute_sliced_w2(samples, name, D, **kw)
    # noise floor: SW2 between two independent ground-truth draws of the same size
    ref = get_reference_sampler(name, D, **kw)
    from mcmc_b200.benchmarks.metrics import sliced_wasserstein_distance
    floor = sliced_wasserstein_distance(ref(1, 50000), ref(2, 50000))
    return d, z, g, sw2, floor


@pytest.mark.parametrize("sampler", ["hmc", "grahmc-constant", "grahmc-tanh", "grahmc-sigmoid", "grahmc-linear",
                                     "grahmc-sine", "nuts", "rwmh"])
def test_gates_standard_normal(gpu_lib, sampler):
    from mcmc_b200.benchmarks import targets as T
    from mcmc_b200.random import normal
    from mcmc_b200.samplers import hmc_run, nuts_run, rahmc_run, rwMH_run, get_friction_schedule
    tgt = T.standard_normal(D)
    q0 = torch.from_numpy(normal(9, (C, D))).cuda()
    S = 400
    if sampler == "hmc":
        out = hmc_run(1, tgt.log_prob_fn, q0, 0.2, 8, S, burn_in=100)   # eps*L ~ pi/2: near-independent draws
    elif sampler.startswith("grahmc"):
        sch = sampler.split("-")[1]
        # gamma = 0.1: the reference's friction does not conserve H, larger amplitudes cost acceptance
        # (tools/quality_probe.py: gamma 0.5 -> accept 0.07-0.35 and split R-hat ~ 1 + (tau-1)/2n > 1.01)
        out = rahmc_run(2, tgt.log_prob_fn, q0, 0.2, 8, 0.1, 0.5 if sch == "tanh" else 2.0, S, burn_in=100,
                        friction_schedule=get_friction_schedule(sch))
    elif sampler == "nuts":
        out = nuts_run(3, tgt.log_prob_fn, q0, 0.5, S, burn_in=100)
    else:
        S = 12000     # tau ~ 60 in 20-D: split R-hat < 1.01 needs ~ 50 tau draws per half chain
        out = rwMH_run(4, tgt.log_prob_fn, q0[:512], S, 2.38 / np.sqrt(D), burn_in=1000)
    samples = out[0]
    assert torch.isfinite(samples).all()
    d, z, g, sw2, floor = _gates(samples, tgt, "standard_normal", S)
    assert d["rhat_max"] < 1.01, d["rhat_max"]
    assert z["pass"], z["reason"]
    assert g["usable"] and g["quality_pass"], (d["ess_bulk_min"], d["ess_tail_min"])
    assert sw2 < 3 * floor + 0.02, (sw2, floor)
    acc = float(out[2].mean())
    assert 0.1 < acc <= 1.0


@pytest.mark.parametrize("name,kw,eps,L,mass", [("correlated_gaussian", {"correlation": 0.9}, 0.12, 48, False),
                                                ("ill_conditioned_gaussian", {"condition_number": 100.0}, 0.2, 8, True)])
def test_gates_gaussians_hmc(gpu_lib, name, kw, eps, L, mass):
    from mcmc_b200.benchmarks import targets as T
    from mcmc_b200.random import normal
    from mcmc_b200.samplers import hmc_run
    tgt = T.get_target(name, D, **kw)
    q0 = torch.from_numpy(normal(11, (C, D))).cuda()
    w = None
    if mass:    # the true variances as diagonal M^-1 (what the warmup learns): the metric path of the kernel
        w = torch.from_numpy(np.diag(tgt.true_cov).copy()).cuda()
        q0 = q0 * w.sqrt()
    samples, lps, acc, st = hmc_run(5, tgt.log_prob_fn, q0, eps, L, 600, burn_in=300, inv_mass_matrix=w)
    d, z, g, sw2, floor = _gates(samples, tgt, name, 600, **kw)
    assert d["rhat_max"] < 1.01 and z["pass"], (d["rhat_max"], z["reason"])
    assert g["usable"]
    assert sw2 < 2 * floor + 0.05, (sw2, floor)
    # sample covariance diagonal against the truth
    var = samples.reshape(-1, D).var(dim=0).cpu().numpy()
    np.testing.assert_allclose(var, np.diag(tgt.true_cov), rtol=0.08)


def test_warmup_on_device_learns_metric_and_matches_oracle_backend(gpu_lib, oracle, monkeypatch):
    """run_adaptive_warmup through the CUDA kernels (strict build: bit-comparable arithmetic on a polynomial
    target) against the same host logic driven by the CPU oracle."""
    import importlib
    from mcmc_b200.benchmarks import targets as T
    from mcmc_b200.random import normal
    from mcmc_b200.tuning import run_adaptive_warmup
    import helpers
    Dw, Cw = 6, 48
    tgt = T.ill_conditioned_gaussian(Dw, condition_number=25.0)
    q0 = normal(21, (Cw, Dw))
    monkeypatch.setenv("MCMC_B200_STRICT", "1")
    eps_g, w_g, pos_g, info_g = run_adaptive_warmup("hmc", tgt.log_prob_fn, None, q0, key=7, num_warmup=2500,
                                                    num_steps=8, verbose=False)
    monkeypatch.delenv("MCMC_B200_STRICT")
    assert w_g.is_cuda and pos_g.is_cuda and info_g["num_da_updates"] == 26
    with monkeypatch.context() as m:
        helpers.patch_tuning_to_oracle(m)
        eps_o, w_o, pos_o, _ = run_adaptive_warmup("hmc", tgt.log_prob_fn, None, q0, key=7, num_warmup=2500,
                                                   num_steps=8, verbose=False)
    assert eps_g == pytest.approx(eps_o, rel=1e-9)
    np.testing.assert_allclose(w_g.cpu().numpy(), w_o.numpy(), rtol=1e-9)
    np.testing.assert_allclose(pos_g.cpu().numpy(), pos_o.numpy(), rtol=1e-7, atol=1e-9)


@pytest.mark.parametrize("sampler,sched", [("hmc", None), ("nuts", None), ("grahmc", "tanh")])
def test_warmup_large_c_fast_path(gpu_lib, sampler, sched):
    from mcmc_b200.benchmarks import targets as T
    from mcmc_b200.random import normal
    from mcmc_b200.tuning import run_adaptive_warmup
    Dw, Cw = 20, 4096
    tgt = T.ill_conditioned_gaussian(Dw, condition_number=100.0)
    q0 = torch.from_numpy(normal(3, (Cw, Dw))).cuda()
    eps, w, pos, info = run_adaptive_warmup(sampler, tgt.log_prob_fn, None, q0, key=5, num_warmup=2500,
                                            schedule_type=sched, num_steps=16, verbose=False, max_iter_step=200)
    true_var = np.linspace(1.0, 100.0, Dw)
    np.testing.assert_allclose(w.cpu().numpy(), 1000 / 1005 * true_var + 5 / 1005, rtol=0.15)   # per-chain window variances are biased low by autocorrelation (same estimator as adaptation.py:240-248)
    assert 0.05 < eps < 3.0
    if sampler == "grahmc":
        assert 0.01 <= info["gamma"] <= 5.0 and info["steepness"] == 0.5 and "friction_tuning" in info


# ---------------------------------------------------------------------------------------------------------------
# The reference's own flow on the remaining BASELINE targets (run_benchmarks.py:468-712: run_adaptive_warmup -> *_run ->
# gates), under the native Philox stream.  Settings probed with the CPU oracle (same flow, 512 chains): student_t and
# log_gamma pass the full quality gate with 2000 draws after the 2500-step warmup; the bimodal mixture passes the hard
# (usable) gate under NUTS.  The Bonferroni z-test has a 5% false-alarm rate by construction, so the assertion uses
# |z| < 4 (0.1%) and reports the literal gate.

@pytest.mark.parametrize("name,sampler,S,gate", [("student_t", "hmc", 2000, "quality"), ("log_gamma", "hmc", 2000, "quality"),
                                                 ("gaussian_mixture", "nuts", 1000, "usable")])
def test_gates_reference_flow_heavy_tailed_positive_and_bimodal_targets(gpu_lib, name, sampler, S, gate):
    from mcmc_b200.benchmarks import targets as T
    from mcmc_b200.samplers import hmc_run, nuts_run
    from mcmc_b200.tuning import run_adaptive_warmup
    tgt = T.get_target(name, D)
    q0 = torch.as_tensor(np.asarray(tgt.init_sampler(3, C)), dtype=torch.float64).cuda()
    eps, w, pos, info = run_adaptive_warmup(sampler, tgt.log_prob_fn, None, q0, key=7, num_warmup=2500, num_steps=16,
                                            verbose=False)
    assert info["num_da_updates"] == 26
    if sampler == "hmc":
        out = hmc_run(1, tgt.log_prob_fn, pos, eps, 16, S, inv_mass_matrix=w)
    else:
        out = nuts_run(1, tgt.log_prob_fn, pos, eps, S, inv_mass_matrix=w)
    samples = out[0]
    assert torch.isfinite(samples).all()
    d, z, g, sw2, floor = _gates(samples, tgt, name, S)
    assert g["usable"], (d["rhat_max"], d["ess_bulk_min"], d["ess_tail_min"])
    assert z["max_z"] < 4.0, z["reason"]
    if gate == "quality":
        assert d["rhat_max"] < 1.01 and d["ess_tail_min"] >= 200, (d["rhat_max"], d["ess_tail_min"])
    assert sw2 < 3 * floor + 0.05, (sw2, floor)
    if name == "log_gamma":
        assert float(samples.min()) > 0.0            # the support: a proposal outside has log p = -inf and is rejected


@pytest.mark.parametrize("name,sampler,L", [("rosenbrock", "hmc", 16), ("rosenbrock", "hmc", 64), ("neals_funnel", "nuts", 0)])
def test_gates_that_fail_fail_identically_on_the_oracle(gpu_lib, oracle, name, sampler, L):
    """BASELINE configs 2 / 5 run the reference's way (warmup -> sample, 256 draws) do NOT pass the reference's hard gate:
    on the 20-D rosenbrock(scale 0.1) split R-hat stays at 1.2-1.4 (the CPU oracle still reports 1.08 after 10,000 draws
    per chain: the ridge mixes slowly under a diagonal metric), and on Neal's funnel the neck is unreachable at one step
    size.  That is a property of the algorithm, not of the kernels: the CPU oracle run through the SAME host flow gives
    the same statistics.  (Different chaotic realisations, same distribution: compared statistically.)"""
    import helpers
    from mcmc_b200 import _abi, diagnostics as DG
    from mcmc_b200.benchmarks import targets as T
    from mcmc_b200.samplers import hmc_run, nuts_run
    from mcmc_b200.tuning import run_adaptive_warmup
    from oracle import diagnostics_np as DN
    Cq, S = 2048, 256
    tgt = T.get_target(name, D)
    tag = T.resolve_log_prob(tgt.log_prob_fn)
    q0 = np.asarray(tgt.init_sampler(3, Cq))
    kw = dict(key=7, num_warmup=2500, num_steps=max(L, 1), verbose=False)
    eps, w, pos, _ = run_adaptive_warmup(sampler, tgt.log_prob_fn, None, torch.from_numpy(q0).cuda(), **kw)
    if sampler == "hmc":
        smp = hmc_run(1, tgt.log_prob_fn, pos, eps, L, S, inv_mass_matrix=w)[0]
    else:
        smp = nuts_run(1, tgt.log_prob_fn, pos, eps, S, inv_mass_matrix=w)[0]
    dg = DG.compute_diagnostics(smp)
    with pytest.MonkeyPatch.context() as m:
        helpers.patch_tuning_to_oracle(m)
        eps_o, w_o, pos_o, _ = run_adaptive_warmup(sampler, tgt.log_prob_fn, None, q0, **kw)
    p = oracle.make_params(tag.target_id, tag.target_params, Cq, D, S, 0, L, eps_o, rng_mode=_abi.RNG_PHILOX, seed=(0, 1))
    sid = _abi.SAMPLER_HMC if sampler == "hmc" else _abi.SAMPLER_NUTS
    do = DN.compute_diagnostics(oracle.run(sid, p, pos_o.numpy(), inv_mass=w_o.numpy())["samples"])
    # two realisations of the same warmup: the learned step size / metric agree statistically (the funnel's variances
    # are set by how deep into the neck a few chains got)
    assert eps == pytest.approx(eps_o, rel=0.1 if name == "rosenbrock" else 0.4)
    np.testing.assert_allclose(w.cpu().numpy(), w_o.numpy(), rtol=0.35 if name == "rosenbrock" else 1.0)
    assert dg["rhat_max"] >= 1.05 and do["rhat_max"] >= 1.05, (dg["rhat_max"], do["rhat_max"])     # usable gate: both fail
    # the funnel's R-hat is set by a handful of chains deep in the neck and moves by +-30% between realisations
    rtol = 0.15 if name == "rosenbrock" else 0.45
    assert abs(dg["rhat_max"] - do["rhat_max"]) < rtol * do["rhat_max"], (dg["rhat_max"], do["rhat_max"])
    assert 0.5 < dg["ess_bulk_min"] / do["ess_bulk_min"] < 2.0, (dg["ess_bulk_min"], do["ess_bulk_min"])
