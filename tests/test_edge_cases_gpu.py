"""Degenerate shapes through the C ABI against the oracle: no stored samples (burn-in only), a single chain, no chains,
a 1-D init_position through the reference-signature wrappers (promoted to one chain, HMC.py:52-54) and the reference's
ValueError for a 3-D one (HMC.py:57)."""
import numpy as np
import pytest

from helpers import TARGET_CASES, GpuBackend, relerr, tid
from mcmc_b200 import _abi

pytestmark = pytest.mark.gpu
SAMPLERS = [(_abi.SAMPLER_HMC, 8), (_abi.SAMPLER_GRAHMC, 8), (_abi.SAMPLER_RWMH, 0), (_abi.SAMPLER_NUTS, 0)]


def _run_both(oracle, sampler, L, C_, S, B, D=20, name="neals_funnel", variant="fast"):
    params, eps, init = TARGET_CASES[name]
    rng = np.random.default_rng(77)
    T_ = B + S
    q0 = init(rng, C_, D)
    kw = dict(inv_mass=np.linspace(0.5, 2.0, D), z=rng.standard_normal((T_, C_, D)), u=rng.random((T_, C_)))
    if sampler == _abi.SAMPLER_NUTS:
        kw.update(dir_u=rng.random((T_, C_, 5)), prop_u=rng.random((T_, C_, 5)))
    if sampler == _abi.SAMPLER_RWMH:
        kw.pop("inv_mass")
    p = oracle.make_params(tid(name), params, C_, D, S, B, L, eps, gamma=1.0, steepness=0.5, schedule_id=1,
                           rng_mode=_abi.RNG_INJECTED, max_tree_depth=5)
    return GpuBackend(variant).run(sampler, p, q0, **kw), oracle.run(sampler, p, q0, **kw)


@pytest.mark.parametrize("sampler,L", SAMPLERS)
def test_burn_in_only_run_stores_nothing_and_advances_the_state(gpu_lib, oracle, sampler, L):
    r, o = _run_both(oracle, sampler, L, C_=37, S=0, B=4)
    assert r["samples"].shape == (0, 37, 20) and r["log_probs"].shape == (0, 37)
    assert relerr(r["final_q"], o["final_q"]) <= 1e-10 and relerr(r["final_lp"], o["final_lp"]) <= 1e-10
    assert np.array_equal(r["accept_count"], o["accept_count"])   # counters reset after burn-in (HMC.py:276-282): zeros
    assert np.all(r["accept_count"] == 0)
    assert np.all(np.isnan(r["accept_rate"]))                      # 0 / 0 in float32, as count / num_samples gives in jax


@pytest.mark.parametrize("sampler,L", SAMPLERS)
def test_empty_run_returns_the_init_state(gpu_lib, oracle, sampler, L):
    """burn_in = num_samples = 0: the reference returns the *_init state (value [+ gradient] at the initial position,
    zero counters, accept_rate = 0 / 0) — HMC.py:259-314, NUTS.py:479-514, RWMH.py:148-178."""
    r, o = _run_both(oracle, sampler, L, C_=37, S=0, B=0)
    q0 = TARGET_CASES["neals_funnel"][2](np.random.default_rng(77), 37, 20)
    assert np.array_equal(r["final_q"], q0)
    assert relerr(r["final_lp"], o["final_lp"]) <= 1e-10
    if sampler != _abi.SAMPLER_RWMH:
        assert relerr(r["final_g"], o["final_g"]) <= 1e-10
        assert np.any(r["final_g"] != 0.0)
    assert np.all(r["accept_count"] == 0) and np.all(np.isnan(r["accept_rate"]))
    assert r["samples"].shape == (0, 37, 20)


@pytest.mark.parametrize("sampler,L", SAMPLERS)
def test_single_chain(gpu_lib, oracle, sampler, L):
    r, o = _run_both(oracle, sampler, L, C_=1, S=5, B=2)
    for k in ("samples", "log_probs", "final_q", "final_lp"):
        assert relerr(r[k], o[k]) <= 1e-10, k
    assert np.array_equal(r["accept_count"], o["accept_count"])


@pytest.mark.parametrize("sampler,L", SAMPLERS)
def test_no_chains_is_a_no_op(gpu_lib, oracle, sampler, L):
    r, _ = _run_both(oracle, sampler, L, C_=0, S=3, B=1)
    assert r["samples"].shape == (3, 0, 20) and r["final_q"].shape == (0, 20)
    assert int(r["counters"][0]) == 0


def test_wrappers_promote_1d_init_and_reject_3d(gpu_lib):
    import torch
    from mcmc_b200 import samplers as SM
    from mcmc_b200.benchmarks import targets as T
    f = T.standard_normal(dim=10).log_prob_fn
    x = torch.linspace(-1, 1, 10, dtype=torch.float64)
    s, lp, ar, st = SM.hmc_run(0, f, x, 0.1, 4, 6, burn_in=2)
    assert s.shape == (6, 1, 10) and lp.shape == (6, 1) and ar.shape == (1,) and ar.dtype == torch.float32
    assert st.position.shape == (1, 10) and st.accept_count.dtype == torch.int32
    s, lp, ar, st, d, m = SM.nuts_run(0, f, x, 0.3, 3)
    assert s.shape == (3, 1, 10) and d.shape == (3, 1) and m.shape == (3, 1)
    for fn, args in ((SM.hmc_run, (0.1, 4, 3)), (SM.rwMH_run, (3, 0.5))):
        with pytest.raises(ValueError):
            fn(0, f, torch.zeros(2, 3, 10, dtype=torch.float64), *args)
    with pytest.raises(KeyError):
        SM.get_friction_schedule("no_such_schedule")   # GRAHMC.py:122
    with pytest.raises(NotImplementedError):
        SM.hmc_run(0, lambda v: -0.5 * (v ** 2).sum(), x, 0.1, 4, 3)   # untagged callable: no CPU fallback


@pytest.mark.parametrize("sampler,L", SAMPLERS)
def test_resume_from_final_state_position_equals_one_run(gpu_lib, oracle, sampler, L):
    """Sampler-level resume as the reference does it (adaptation.py:215, run_benchmarks.py:614): the next call starts
    from final_state.position and re-derives log-prob / gradient.  With the same draws, 4 + 5 iterations in two calls
    reproduce 9 iterations in one call bit for bit."""
    name, D, C_ = "rosenbrock", 20, 70
    params, eps, init = TARGET_CASES[name]
    rng = np.random.default_rng(5)
    q0 = init(rng, C_, D)
    z, u = rng.standard_normal((9, C_, D)), rng.random((9, C_))
    du, pu = rng.random((9, C_, 5)), rng.random((9, C_, 5))

    def run(q, lo, hi):
        kw = dict(z=z[lo:hi], u=u[lo:hi])
        if sampler == _abi.SAMPLER_NUTS:
            kw.update(dir_u=du[lo:hi], prop_u=pu[lo:hi])
        p = oracle.make_params(tid(name), params, C_, D, hi - lo, 0, L, eps, gamma=1.0, steepness=0.5, schedule_id=1,
                               rng_mode=_abi.RNG_INJECTED, max_tree_depth=5)
        return GpuBackend("fast").run(sampler, p, q, **kw)

    whole = run(q0, 0, 9)
    first = run(q0, 0, 4)
    second = run(first["final_q"], 4, 9)
    assert np.array_equal(np.concatenate([first["samples"], second["samples"]]), whole["samples"])
    assert np.array_equal(np.concatenate([first["log_probs"], second["log_probs"]]), whole["log_probs"])
    assert np.array_equal(first["accept_count"] + second["accept_count"], whole["accept_count"])
