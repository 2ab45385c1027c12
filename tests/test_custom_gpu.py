"""User-supplied densities on the GPU (mcmc_b200/custom.py): a user-written rosenbrock / normal, compiled into the
sampler kernels at run time, against the built-in targets with the same injected draws — all four samplers, values at
1e-10 (the summation order of the user's loop differs from the built-in code), accept counts and tree depths identical."""
import numpy as np
import pytest

from custom_cases import ROSENBROCK, SCALED_NORMAL
from helpers import GpuBackend, relerr, tid
from mcmc_b200 import _abi, custom

pytestmark = pytest.mark.gpu
SAMPLERS = [(_abi.SAMPLER_HMC, 12), (_abi.SAMPLER_GRAHMC, 12), (_abi.SAMPLER_RWMH, 0), (_abi.SAMPLER_NUTS, 0)]


@pytest.fixture(scope="module")
def user_targets(tmp_path_factory):
    import os
    os.environ["MCMC_B200_CACHE"] = str(tmp_path_factory.mktemp("user_targets"))
    return {"rosenbrock": custom.compile_target("my_rosenbrock", ROSENBROCK, dim=20, params=[100.0]),
            "normal": custom.compile_target("my_normal", SCALED_NORMAL, dim=20, params=[1.0])}


@pytest.mark.parametrize("D", [20, 7])
@pytest.mark.parametrize("sampler,L", SAMPLERS)
@pytest.mark.parametrize("which", ["rosenbrock", "normal"])
def test_user_target_matches_the_builtin_one(gpu_lib, oracle, user_targets, which, sampler, L, D):
    ut = user_targets[which].log_prob_fn
    builtin, bpar, upar, eps = (("rosenbrock", [0.1], [100.0], 0.005) if which == "rosenbrock"
                                else ("standard_normal", [0.0], [1.0], 0.2))
    C_, S, B, md = 90, 5, 2, 6
    rng = np.random.default_rng(3)
    q0 = (1.0 + 0.05 * rng.standard_normal((C_, D))) if which == "rosenbrock" else rng.standard_normal((C_, D))
    kw = dict(z=rng.standard_normal((B + S, C_, D)), u=rng.random((B + S, C_)))
    if sampler == _abi.SAMPLER_NUTS:
        kw.update(dir_u=rng.random((B + S, C_, md)), prop_u=rng.random((B + S, C_, md)))
    if sampler != _abi.SAMPLER_RWMH:
        kw["inv_mass"] = np.linspace(0.5, 2.0, D)
    step = 0.3 if sampler == _abi.SAMPLER_RWMH and which == "normal" else (0.01 if sampler == _abi.SAMPLER_RWMH else eps)
    mk = lambda t, par: oracle.make_params(t, par, C_, D, S, B, L, step, gamma=0.5, steepness=0.5, schedule_id=1,
                                           rng_mode=_abi.RNG_INJECTED, max_tree_depth=md)
    g = GpuBackend("fast")
    ref = g.run(sampler, mk(tid(builtin), bpar), q0, **kw)
    got = g.run(sampler, mk(ut.target_id, upar), q0, **kw)
    keys = ["samples", "log_probs", "final_q", "final_lp"] + ([] if sampler == _abi.SAMPLER_RWMH else ["final_g"])
    if sampler == _abi.SAMPLER_NUTS:
        keys.append("mean_accept_probs")
        assert np.array_equal(got["tree_depths"], ref["tree_depths"])
    for k in keys:
        assert relerr(got[k], ref[k]) <= 1e-10, (k, relerr(got[k], ref[k]))
    assert np.array_equal(got["accept_count"], ref["accept_count"])


def test_user_target_through_the_reference_signatures(gpu_lib, user_targets):
    """hmc_run / nuts_run with the tagged closure, native Philox: moments of N(0, 1/2) from par[0] = 2."""
    import torch
    from mcmc_b200.samplers import hmc_run, nuts_run
    t = custom.compile_target("my_normal", SCALED_NORMAL, dim=20, params=[2.0])
    q0 = torch.randn((2048, 20), dtype=torch.float64, device="cuda")
    smp, lp, acc, st = hmc_run(1, t.log_prob_fn, q0, 0.15, 8, 300, burn_in=100)
    assert abs(float(smp.var()) - 0.5) < 0.02 and abs(float(smp.mean())) < 0.01 and 0.5 < float(acc.mean()) <= 1.0
    lp2 = t.log_prob_fn(smp[-1])
    assert torch.allclose(lp2, lp[-1], rtol=1e-12)
    out = nuts_run(2, t.log_prob_fn, q0, 0.3, 100, burn_in=50)
    assert abs(float(out[0].var()) - 0.5) < 0.03
