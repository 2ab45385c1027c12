"""The CUDA path (through the C ABI, both builds) against outputs of the REFERENCE'S OWN SOURCE
(tests/golden/reference_golden.npz — HMC.py / GRAHMC.py / RWMH.py / NUTS.py / targets.py executed unmodified on a
torch-fp64 stand-in for jax, see tests/golden/make_reference_golden.py) with the reference run's own momenta and
uniforms injected: every stored position, log-prob, proposal, dH and final gradient within 1e-10 relative
(|a-b| / max(|b|, 1)), every accept decision and NUTS tree depth identical.  No oracle in this comparison."""
import numpy as np
import pytest

from helpers import GpuBackend, relerr
from reference_cases import EXACT_KEYS, FLOAT_KEYS, load, run_case

pytestmark = pytest.mark.gpu
CASES = load()
TOL = 1e-10


@pytest.mark.parametrize("variant", ["strict", "fast"])
@pytest.mark.parametrize("cid", sorted(CASES))
def test_cuda_matches_reference_source(gpu_lib, cid, variant):
    c = CASES[cid]
    got = run_case(GpuBackend(variant), c)
    for k in EXACT_KEYS:
        if k in c:
            assert np.array_equal(got[k].reshape(c[k].shape), c[k]), (cid, k)
    for k in FLOAT_KEYS:
        if k in c and k in got:
            e = relerr(got[k].reshape(c[k].shape), c[k])
            assert e <= TOL, (cid, variant, k, e)
    assert np.array_equal(got["accept_rate"], c["accept_rate"]), cid


def test_quickstart_signature_matches_reference_source(gpu_lib):
    """The torch device-pointer runtime under the reference-signature wrappers (mcmc_b200.runtime.run_sampler, the
    layer mcmc_b200.samplers.rahmc_run calls) on the fixture's D=10 GRAHMC case, friction_schedule=None -> constant."""
    import torch
    from mcmc_b200 import _abi
    from mcmc_b200.benchmarks import targets as T
    from mcmc_b200.runtime import run_sampler
    c = CASES["grahmc/none/standard_normal_D10"]
    m = c["meta"]
    tag = T.standard_normal(dim=m["D"]).log_prob_fn
    out = run_sampler(_abi.SAMPLER_GRAHMC, tag, torch.from_numpy(c["q0"]).cuda(), key=0, step_size=float(c["eps"]),
                      num_steps=m["L"], num_samples=m["S"], burn_in=m["B"], gamma=m["gamma"], steepness=1.0, schedule_id=0,
                      inv_mass_matrix=torch.from_numpy(c["inv_mass"]).cuda(),
                      inject=dict(z=torch.from_numpy(c["z"]).cuda(), u=torch.from_numpy(c["u"]).cuda()))
    assert relerr(out["samples"].cpu().numpy(), c["samples"]) <= TOL
    assert relerr(out["log_probs"].cpu().numpy(), c["log_probs"]) <= TOL
