"""dense_gaussian (target id 8) on the GPU: value + gradient against the reference-source fixtures of
correlated_gaussian (the same `x @ cov_inv @ x`, targets.py:94) and every sampler against the oracle with injected
draws, both builds, compiled and runtime dimensions; native-stream gates are in tests/test_quality_gpu.py."""
import numpy as np
import pytest
import torch

from dense_cases import compound_symmetry_cov, dense_target, random_spd
from helpers import GpuBackend, relerr
from mcmc_b200 import _abi
from mcmc_b200.benchmarks import targets as T
from mcmc_b200.runtime import target_eval
from test_reference_host_golden_cpu import BLOB, TARGET_KEYS, parse_target_key

pytestmark = pytest.mark.gpu
CORR_KEYS = [k for k in TARGET_KEYS if "|correlated_gaussian|" in k]
TID = _abi.TARGET_IDS["dense_gaussian"]


@pytest.mark.parametrize("variant", ["strict", "fast"])
@pytest.mark.parametrize("key", CORR_KEYS)
def test_cuda_dense_gaussian_equals_reference_correlated_gaussian(gpu_lib, key, variant):
    _, D, kw = parse_target_key(key)
    tag = T.resolve_log_prob(dense_target(compound_symmetry_cov(D, kw["correlation"])).log_prob_fn)
    lp, g = target_eval(tag, torch.from_numpy(BLOB[key + "|x"]).cuda(), variant=variant)
    assert relerr(lp.cpu().numpy(), BLOB[key + "|lp"]) <= 1e-12
    assert relerr(g.cpu().numpy(), BLOB[key + "|grad"]) <= 1e-11


def _case(D, C_, S, B, md, seed):
    cov = random_spd(D, seed=seed)
    tag = T.resolve_log_prob(dense_target(cov).log_prob_fn)
    rng = np.random.default_rng(seed)
    T_ = B + S
    chol = np.linalg.cholesky(cov)
    return dict(tag=tag, q0=rng.standard_normal((C_, D)) @ chol.T, z=rng.standard_normal((T_, C_, D)),
                u=rng.random((T_, C_)), du=rng.random((T_, C_, md)), pu=rng.random((T_, C_, md)),
                w=np.linspace(0.5, 2.0, D))


@pytest.mark.parametrize("variant", ["strict", "fast"])
@pytest.mark.parametrize("D", [20, 10, 7])
@pytest.mark.parametrize("sampler", [_abi.SAMPLER_HMC, _abi.SAMPLER_GRAHMC, _abi.SAMPLER_RWMH, _abi.SAMPLER_NUTS])
def test_dense_gaussian_sampler_parity(gpu_lib, oracle, sampler, D, variant):
    C_, S, B, md = 100, 5, 2, 6
    c = _case(D, C_, S, B, md, seed=40 + D)
    hmc = sampler in (_abi.SAMPLER_HMC, _abi.SAMPLER_GRAHMC)
    eps = 0.4 if sampler == _abi.SAMPLER_RWMH else 0.04
    p = oracle.make_params(TID, c["tag"].target_params, C_, D, S, B, 16 if hmc else 0, eps, gamma=1.0, steepness=0.5,
                           schedule_id=1, max_tree_depth=md, flags=_abi.FLAG_TRACK_PROPOSALS if hmc else 0)
    kw = dict(cov_inv=c["tag"].cov_inv, z=c["z"], u=c["u"])
    if sampler != _abi.SAMPLER_RWMH:
        kw["inv_mass"] = c["w"]
    if sampler == _abi.SAMPLER_NUTS:
        kw.update(dir_u=c["du"], prop_u=c["pu"])
    o = oracle.run(sampler, p, c["q0"], **kw)
    r = GpuBackend(variant).run(sampler, p, c["q0"], **kw)
    tol = 1e-12 if variant == "strict" else 1e-10
    keys = ["samples", "log_probs", "final_q", "final_lp"]
    if sampler != _abi.SAMPLER_RWMH:
        keys.append("final_g")
    if hmc:
        keys += ["prop_q", "prop_lp", "delta_h", "pre_q", "pre_lp"]
    if sampler == _abi.SAMPLER_NUTS:
        keys.append("mean_accept_probs")
        assert np.array_equal(r["tree_depths"], o["tree_depths"])
    for k in keys:
        assert relerr(r[k], o[k]) <= tol, (k, relerr(r[k], o[k]))
    assert np.array_equal(r["accept_count"], o["accept_count"])
    assert int(o["accept_count"].sum()) > 0


def test_dense_gaussian_wrapper_errors(gpu_lib):
    """The C ABI refuses the target without its matrix; the wrapper passes it from the tagged closure."""
    from mcmc_b200.hostrun import make_params, run_host
    q0 = np.zeros((4, 20))
    with pytest.raises(ValueError):
        run_host(_abi.SAMPLER_HMC, make_params(TID, [0.0], 4, 20, 2, num_steps=2, step_size=0.1), q0)
    from mcmc_b200.samplers import hmc_run
    t = dense_target(random_spd(20, 3))
    s, lp, ar, st = hmc_run(1, t.log_prob_fn, torch.zeros(4, 20, dtype=torch.float64), 0.05, 4, 3)
    assert s.shape == (3, 4, 20) and torch.isfinite(lp).all()
