"""The oracle's stand-alone integrator (oracle_leapfrog) against the REFERENCE'S OWN `leapfrog` (samplers/HMC.py:85-134)
and `_trajectory_with_schedule` (samplers/GRAHMC.py:215-261), run by tests/golden/make_reference_leapfrog_golden.py:
same (position, momentum, log_prob, grad_log_prob) in -> the same four arrays out within 1e-10 relative, for every
benchmark target and every friction schedule; plus the known-answer properties of SURVEY.md section 8(c)."""
import numpy as np
import pytest

from helpers import TARGET_CASES, relerr, tid
from leapfrog_cases import OUT_KEYS, load, run_oracle

CASES = load()
TOL = 1e-10


@pytest.mark.parametrize("cid", sorted(CASES))
def test_oracle_leapfrog_matches_reference_source(oracle, cid):
    c = CASES[cid]
    got = run_oracle(oracle, c)
    for k in OUT_KEYS:
        assert got[k].shape == c[k].shape, (cid, k)
        e = relerr(got[k], c[k])      # relerr treats equal infinities / NaNs in the same place as agreement
        assert e <= TOL, (cid, k, e)


def test_fixture_covers_every_target_and_schedule():
    metas = [c["meta"] for c in CASES.values()]
    assert {m["target"] for m in metas if m["fn"] == "leapfrog"} == set(TARGET_CASES)
    assert {m["sched"] for m in metas if m["fn"] == "trajectory"} == {"constant", "tanh", "sigmoid", "linear", "sine"}
    assert {m["D"] for m in metas} >= {20, 10, 7}            # both compiled dimensions and the runtime-dimension kernels
    assert {m["L"] for m in metas} >= {0, 1}
    # the log-gamma trajectory leaves the support: the value is -inf and the gradient the reference's zeros
    lg = CASES["leapfrog/log_gamma"]
    assert np.isneginf(lg["lp1"]).any()


def test_zero_steps_return_the_inputs(oracle):
    c = CASES["leapfrog/student_t/L0"]
    got = run_oracle(oracle, c)
    for k_out, k_in in (("q1", "q0"), ("p1", "p0"), ("lp1", "lp0"), ("g1", "g0")):
        assert np.array_equal(got[k_out], c[k_in]) and np.array_equal(c[k_out], c[k_in])


@pytest.mark.parametrize("name", sorted(TARGET_CASES))
def test_oracle_leapfrog_is_reversible(oracle, name):
    """Integrate, negate the momentum, integrate again, negate: back at the start (leapfrog is time-reversible)."""
    par, eps, init = TARGET_CASES[name]
    r = np.random.default_rng(7)
    C_, D, L = 16, 20, 24
    q0 = init(r, C_, D)
    w = np.linspace(0.5, 2.0, D)
    p0 = r.standard_normal((C_, D)) / np.sqrt(w)
    eps = eps * 0.25                                        # stays inside the support of log_gamma
    p = oracle.make_params(tid(name), par, C_, D, 0, 0, L, eps)
    lp0, g0 = oracle.target_eval(p, q0)
    q1, p1, lp1, g1 = oracle.leapfrog(p, q0, p0, lp0, g0, inv_mass=w)
    q2, p2, lp2, g2 = oracle.leapfrog(p, q1, -p1, lp1, g1, inv_mass=w)
    assert np.isfinite(lp1).all()
    assert relerr(q2, q0) < 1e-9 and relerr(-p2, p0) < 1e-9 and relerr(lp2, lp0) < 1e-9


def test_zero_friction_trajectory_is_the_plain_leapfrog(oracle):
    """gamma = 0 -> every s_i = exp(0) = 1 exactly: the conformal trajectory equals `leapfrog` bit for bit."""
    par, eps, init = TARGET_CASES["rosenbrock"]
    r = np.random.default_rng(8)
    q0 = init(r, 8, 20)
    p0 = r.standard_normal((8, 20))
    p = oracle.make_params(tid("rosenbrock"), par, 8, 20, 0, 0, 16, eps, gamma=0.0, steepness=0.5, schedule_id=1)
    lp0, g0 = oracle.target_eval(p, q0)
    a = oracle.leapfrog(p, q0, p0, lp0, g0)
    b = oracle.leapfrog(p, q0, p0, lp0, g0, friction=True)
    for x, y in zip(a, b):
        assert np.array_equal(x, y)


def test_oracle_leapfrog_agrees_with_the_sampler_loop(oracle):
    """The proposal of one HMC iteration (tracked prop_q / prop_lp) is the stand-alone trajectory from the same state
    and momentum: the two restatements of HMC.py:86-134 inside the oracle are the same arithmetic."""
    from mcmc_b200 import _abi
    par, eps, init = TARGET_CASES["neals_funnel"]
    r = np.random.default_rng(9)
    C_, D, L = 6, 20, 10
    q0 = init(r, C_, D)
    w = np.linspace(0.5, 2.0, D)
    z = r.standard_normal((1, C_, D))
    u = np.full((1, C_), 0.5)
    p = oracle.make_params(tid("neals_funnel"), par, C_, D, 1, 0, L, eps, flags=_abi.FLAG_TRACK_PROPOSALS)
    out = oracle.run(_abi.SAMPLER_HMC, p, q0, inv_mass=w, z=z, u=u)
    lp0, g0 = oracle.target_eval(p, q0)
    q1, p1, lp1, g1 = oracle.leapfrog(p, q0, z[0] / np.sqrt(w), lp0, g0, inv_mass=w)
    assert np.array_equal(out["prop_q"][0], q1) and np.array_equal(out["prop_lp"][0], lp1)
