"""mcmc_leapfrog (csrc/leapfrog_kernel.cuh) — the trajectory as a stand-alone launch — through the reference-named Python
functions `samplers.HMC.leapfrog` (HMC.py:85-134) and `samplers.GRAHMC._trajectory_with_schedule` (GRAHMC.py:215-261):
 * against the REFERENCE'S OWN SOURCE on the recorded inputs (tests/golden/reference_leapfrog_golden.npz), both builds,
   1e-10 relative, every target and every friction schedule, D = 20 / 10 / 7 (runtime-dimension kernels);
 * against the C oracle on larger seeded inputs: the strict build BIT-EXACT on the targets without a libm call;
 * at BASELINE config 2's full size (65,536 chains x D = 20, L = 128) by reversibility, and against the proposal the
   sampler kernel itself makes from the same state and momentum (hmc_run with injected draws);
 * edge cases: no chains, zero steps, outputs aliasing inputs, bad arguments."""
import ctypes as C

import numpy as np
import pytest

from helpers import POLYNOMIAL, TARGET_CASES, relerr, tid
from leapfrog_cases import OUT_KEYS, load, run_cuda

pytestmark = pytest.mark.gpu
CASES = load()
TOL = 1e-10


@pytest.mark.parametrize("variant", ["strict", "fast"])
@pytest.mark.parametrize("cid", sorted(CASES))
def test_cuda_leapfrog_matches_reference_source(gpu_lib, cid, variant):
    c = CASES[cid]
    got = run_cuda(c, variant)
    for k in OUT_KEYS:
        assert got[k].shape == c[k].shape, (cid, k)
        e = relerr(got[k], c[k])
        assert e <= TOL, (cid, variant, k, e)


def _inputs(oracle, name, C_, D, L, seed, eps_scale=1.0, **pk):
    par, eps, init = TARGET_CASES[name]
    r = np.random.default_rng(seed)
    q0 = init(r, C_, D)
    w = np.linspace(0.5, 2.0, D)
    p0 = r.standard_normal((C_, D)) / np.sqrt(w)
    p = oracle.make_params(tid(name), par, C_, D, 0, 0, L, eps * eps_scale, **pk)
    lp0, g0 = oracle.target_eval(p, q0)
    return p, q0, p0, lp0, g0, w


def _cuda(p, q0, p0, lp0, g0, w, friction, variant):
    import torch
    from mcmc_b200 import _abi
    from mcmc_b200.benchmarks.targets import TaggedLogProb
    from mcmc_b200.runtime import leapfrog
    names = {v: k for k, v in _abi.TARGET_IDS.items()}
    tag = TaggedLogProb(names[p.target_id], p.n_dim, list(p.target_params))
    out = leapfrog(tag, q0, p0, lp0, g0, p.step_size, p.num_steps, w, friction=friction, gamma=p.gamma, steepness=p.steepness,
                   schedule_id=p.schedule_id, variant=variant)
    torch.cuda.synchronize()
    return [x.cpu().numpy() for x in out]


@pytest.mark.parametrize("friction", [False, True])
@pytest.mark.parametrize("D", [20, 10, 13])
@pytest.mark.parametrize("name", sorted(TARGET_CASES))
def test_cuda_leapfrog_matches_oracle(gpu_lib, oracle, name, D, friction):
    """1,000 chains x 20 steps (ragged against the 128-thread blocks).  Strict build: bit-exact where no libm call is on
    the path (polynomial targets, no friction table), 1e-12 otherwise; fast build 1e-10."""
    p, q0, p0, lp0, g0, w = _inputs(oracle, name, 1000, D, 20, 31, eps_scale=0.5, gamma=0.8, steepness=0.5, schedule_id=1)
    ref = oracle.leapfrog(p, q0, p0, lp0, g0, inv_mass=w, friction=friction)
    strict = _cuda(p, q0, p0, lp0, g0, w, friction, "strict")
    fast = _cuda(p, q0, p0, lp0, g0, w, friction, "fast")
    for a, b, f in zip(ref, strict, fast):
        if name in POLYNOMIAL and not friction:
            assert np.array_equal(a, b), name
        else:
            assert relerr(b, a) <= 1e-12, (name, relerr(b, a))
        assert relerr(f, a) <= 1e-10, (name, relerr(f, a))


def test_fullsize_reversibility(gpu_lib, oracle):
    """BASELINE config 2 size: 65,536 chains x D = 20, L = 128 on the rosenbrock.  Integrate, flip the momentum, integrate
    again: back at the start.  The round-trip error is rounding amplified by the trajectory's sensitivity (the same
    quantity on a 2,048-chain sample of the oracle sets the yardstick), and the energy error stays that of a stable
    leapfrog trajectory."""
    from mcmc_b200.runtime import leapfrog
    from mcmc_b200.benchmarks import targets as T
    import torch
    C_, D, L, eps = 65536, 20, 128, 0.01
    p, q0, p0, lp0, g0, w = _inputs(oracle, "rosenbrock", C_, D, L, 41, eps_scale=eps / TARGET_CASES["rosenbrock"][1])
    tag = T.rosenbrock(dim=D).log_prob_fn
    q1, p1, lp1, g1 = leapfrog(tag, q0, p0, lp0, g0, eps, L, w)
    q2, p2, lp2, g2 = leapfrog(tag, q1, -p1, lp1, g1, eps, L, w)
    back = float((q2.cpu() - torch.from_numpy(q0)).abs().max())
    backp = float((-p2.cpu() - torch.from_numpy(p0)).abs().max())
    n = 2048
    po = oracle.make_params(p.target_id, list(p.target_params), n, D, 0, 0, L, eps)
    a = oracle.leapfrog(po, q0[:n], p0[:n], lp0[:n], g0[:n], inv_mass=w)
    b = oracle.leapfrog(po, a[0], -a[1], a[2], a[3], inv_mass=w)
    yard = max(np.abs(b[0] - q0[:n]).max(), np.abs(-b[1] - p0[:n]).max(), 1e-13)
    assert back <= 50 * yard and backp <= 50 * yard, (back, backp, yard)
    assert back < 1e-7
    # the first 2,048 chains agree with the oracle's forward trajectory
    assert relerr(q1[:n].cpu().numpy(), a[0]) <= 1e-9 and relerr(lp1[:n].cpu().numpy(), a[2]) <= 1e-9
    # energy error of the trajectory (kinetic with the diagonal inverse mass, HMC.py:176)
    wt = torch.from_numpy(w).cuda()
    h0 = -torch.from_numpy(lp0).cuda() + 0.5 * ((torch.from_numpy(p0).cuda() ** 2) * wt).sum(1)
    h1 = -lp1 + 0.5 * ((p1 ** 2) * wt).sum(1)
    assert float((h1 - h0).abs().median()) < 0.5


@pytest.mark.parametrize("variant", ["strict", "fast"])
@pytest.mark.parametrize("friction", [False, True])
def test_fullsize_leapfrog_is_the_sampler_kernels_proposal(gpu_lib, oracle, variant, friction):
    """65,536 chains: the proposal hmc_run / rahmc_run make in their first iteration (tracked prop_q, prop_lp) from the
    injected momenta is the stand-alone trajectory from (init, z / sqrt(w)) — bit for bit in the strict build (same
    operation order), to rounding in the fast build (merged half-kicks and folded eps*w in the sampler kernel)."""
    import torch
    from mcmc_b200 import _abi
    from mcmc_b200.benchmarks import targets as T
    from mcmc_b200.runtime import leapfrog, run_sampler, target_eval
    C_, D, L, eps = 65536, 20, 16, 0.01
    r = np.random.default_rng(43)
    q0 = torch.from_numpy(1.0 + 0.05 * r.standard_normal((C_, D))).cuda()
    w = torch.from_numpy(np.linspace(0.5, 2.0, D)).cuda()
    z = torch.from_numpy(r.standard_normal((1, C_, D))).cuda()
    u = torch.full((1, C_), 0.5, dtype=torch.float64, device="cuda")
    tag = T.rosenbrock(dim=D).log_prob_fn
    kw = dict(gamma=0.7, steepness=0.5, schedule_id=_abi.SCHEDULE_IDS["tanh"]) if friction else {}
    out = run_sampler(_abi.SAMPLER_GRAHMC if friction else _abi.SAMPLER_HMC, tag, q0, key=0, step_size=eps, num_steps=L,
                      num_samples=1, burn_in=0, inv_mass_matrix=w, track_proposals=True, inject=dict(z=z, u=u), variant=variant, **kw)
    lp0, g0 = target_eval(tag, q0, variant=variant)
    q1, p1, lp1, g1 = leapfrog(tag, q0, z[0] / torch.sqrt(w), lp0, g0, eps, L, w, friction=friction, variant=variant, **kw)
    if variant == "strict":
        assert torch.equal(out["prop_q"][0], q1) and torch.equal(out["prop_lp"][0], lp1)
    else:
        assert relerr(out["prop_q"][0].cpu().numpy(), q1.cpu().numpy()) <= 1e-10
        assert relerr(out["prop_lp"][0].cpu().numpy(), lp1.cpu().numpy()) <= 1e-9


def test_conformal_step_and_helper_names(gpu_lib, oracle):
    """`_conformal_leapfrog_step` (GRAHMC.py:152-211) with an explicit gamma is a one-step trajectory; the module-level
    `standard_normal_log_prob` (GRAHMC.py:42-46) works for any dimension; `constant_schedule_default` (:108-110)."""
    import torch
    from mcmc_b200.samplers import GRAHMC
    for D in (20, 6):
        p, q0, p0, lp0, g0, w = _inputs(oracle, "standard_normal", 300, D, 1, 51)
        q, mom, lp, g = GRAHMC._conformal_leapfrog_step(q0, p0, p.step_size, 0.9, lp0, g0, GRAHMC.standard_normal_log_prob, w)
        s = np.exp(-0.9 * 0.5 * p.step_size)
        pe = p0 * s + 0.5 * p.step_size * g0
        qe = q0 + p.step_size * (pe * w)
        pe = (pe + 0.5 * p.step_size * (-qe)) * s
        assert relerr(q.cpu().numpy(), qe) <= 1e-13 and relerr(mom.cpu().numpy(), pe) <= 1e-13
        assert relerr(lp.cpu().numpy(), -0.5 * ((qe ** 2).sum(1) + D * np.log(2 * np.pi))) <= 1e-13
        assert torch.allclose(GRAHMC.standard_normal_log_prob(torch.from_numpy(q0).cuda()).cpu(), torch.from_numpy(lp0), rtol=1e-14)
    assert GRAHMC.constant_schedule_default(0.1, 1.0) == -1.0 and GRAHMC.constant_schedule_default(0.6, 1.0) == 1.0


def test_edge_cases_and_errors(gpu_lib, oracle):
    import torch
    from mcmc_b200 import _abi
    from mcmc_b200.benchmarks import targets as T
    from mcmc_b200.runtime import leapfrog, make_params
    from mcmc_b200.samplers import GRAHMC, HMC
    tgt = T.rosenbrock(dim=20)
    f = tgt.log_prob_fn
    # no chains
    e = torch.empty((0, 20), dtype=torch.float64, device="cuda")
    out = HMC.leapfrog(e, e, 0.01, torch.empty((0,), dtype=torch.float64, device="cuda"), e, f, 5, None)
    assert out[0].shape == (0, 20) and out[3].shape == (0,)
    # zero steps: the inputs come back (scan of length 0), inverse mass None = ones
    p, q0, p0, lp0, g0, w = _inputs(oracle, "rosenbrock", 130, 20, 0, 61)
    q, mom, g, lp = HMC.leapfrog(q0, p0, 0.01, lp0, g0, f, 0, None)
    assert np.array_equal(q.cpu().numpy(), q0) and np.array_equal(mom.cpu().numpy(), p0)
    assert np.array_equal(g.cpu().numpy(), g0) and np.array_equal(lp.cpu().numpy(), lp0)
    # a 1-D position is one chain (HMC.py:43-57)
    q, mom, g, lp = HMC.leapfrog(q0[0], p0[0], 0.01, lp0[:1], g0[0], f, 3, None)
    assert q.shape == (1, 20)
    # outputs aliasing the inputs through the C ABI: same result as separate outputs
    L = _abi.lib("fast")
    dq, dp_, dg, dl = (torch.from_numpy(x).cuda() for x in (q0, p0, g0, lp0))
    ref = leapfrog(f, dq, dp_, dl, dg, 0.01, 9, None)
    pp = make_params(f.target_id, f.target_params, 130, 20, 0, 0, 9, 0.01)
    ptr = lambda t: C.c_void_p(t.data_ptr())
    rc = L.mcmc_leapfrog(C.byref(pp), 0, None, None, ptr(dq), ptr(dp_), ptr(dl), ptr(dg), ptr(dq), ptr(dp_), ptr(dl), ptr(dg),
                         C.c_void_p(torch.cuda.current_stream().cuda_stream))
    torch.cuda.synchronize()
    assert rc == 0 and torch.equal(dq, ref[0]) and torch.equal(dp_, ref[1]) and torch.equal(dl, ref[2]) and torch.equal(dg, ref[3])
    # errors: unknown schedule (KeyError like GRAHMC.py:122), too many steps, missing inputs, wrong dimension
    with pytest.raises(KeyError):
        GRAHMC._trajectory_with_schedule(q0, p0, 0.01, 1.0, 0.5, lp0, g0, 4, f, "no_such_schedule", None)
    with pytest.raises(_abi.McmcError):
        HMC.leapfrog(q0, p0, 0.01, lp0, g0, f, 5000, None)
    pp.schedule_id = 99
    assert L.mcmc_leapfrog(C.byref(pp), 1, None, None, ptr(dq), ptr(dp_), ptr(dl), ptr(dg), None, None, None, None, None) == _abi.EINVAL
    pp.schedule_id = 0
    assert L.mcmc_leapfrog(C.byref(pp), 0, None, None, ptr(dq), None, ptr(dl), ptr(dg), None, None, None, None, None) == _abi.EINVAL
    assert b"required" in L.mcmc_last_error()
    with pytest.raises(ValueError):
        HMC.leapfrog(q0[:, :10], p0[:, :10], 0.01, lp0, g0[:, :10], f, 2, None)


def test_user_target_leapfrog(gpu_lib, oracle, tmp_path_factory):
    """A density compiled at run time (mcmc_b200/custom.py) gets the stand-alone trajectory too."""
    import os
    from custom_cases import ROSENBROCK
    from mcmc_b200 import custom
    from mcmc_b200.samplers import HMC
    from mcmc_b200.benchmarks import targets as T
    os.environ["MCMC_B200_CACHE"] = str(tmp_path_factory.mktemp("user_targets_lf"))
    ut = custom.compile_target("my_rosenbrock_lf", ROSENBROCK, dim=20, params=[100.0])
    p, q0, p0, lp0, g0, w = _inputs(oracle, "rosenbrock", 257, 20, 11, 71)
    a = HMC.leapfrog(q0, p0, p.step_size, lp0, g0, ut.log_prob_fn, 11, w)
    b = HMC.leapfrog(q0, p0, p.step_size, lp0, g0, T.rosenbrock(dim=20).log_prob_fn, 11, w)
    for x, y in zip(a, b):
        assert relerr(x.cpu().numpy(), y.cpu().numpy()) <= 1e-10
