"""CUDA target evaluation and friction tables against the reference's own source
(tests/golden/reference_host_golden.npz): value within 1e-12, analytic gradient vs the reference's autodiff gradient
within 1e-11, gamma(t) and exp(-gamma*eps/2) within 1e-14 relative; both builds."""
import numpy as np
import pytest
import torch

from helpers import relerr
from mcmc_b200 import _abi
from mcmc_b200.benchmarks import targets as T
from mcmc_b200.runtime import friction_table, target_eval
from test_reference_host_golden_cpu import BLOB, FRICTION_KEYS, TARGET_KEYS, parse_target_key

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("variant", ["strict", "fast"])
@pytest.mark.parametrize("key", TARGET_KEYS)
def test_cuda_target_value_and_gradient_match_reference_autodiff(gpu_lib, key, variant):
    name, D, kw = parse_target_key(key)
    tag = T.resolve_log_prob(getattr(T, name)(dim=D, **kw).log_prob_fn)
    lp, g = target_eval(tag, torch.from_numpy(BLOB[key + "|x"]).cuda(), variant=variant)
    assert relerr(lp.cpu().numpy(), BLOB[key + "|lp"]) <= 1e-12
    assert relerr(g.cpu().numpy(), BLOB[key + "|grad"]) <= 1e-11


@pytest.mark.parametrize("variant", ["strict", "fast"])
@pytest.mark.parametrize("key", FRICTION_KEYS)
def test_cuda_friction_table_matches_reference_source(gpu_lib, key, variant):
    _, name, L, eps, gam, steep = key.split("|")
    s, g = friction_table(_abi.SCHEDULE_IDS[name], int(L), float(eps), float(gam), float(steep), device=0, variant=variant)
    want = BLOB[key]
    assert relerr(g.cpu().numpy(), want) <= 1e-14
    assert relerr(s.cpu().numpy(), np.exp(-want * (0.5 * float(eps)))) <= 1e-14


@pytest.mark.parametrize("tag", ["equal", "unequal"])
def test_sliced_w2_matches_reference_source(gpu_lib, tag):
    """benchmarks/metrics.py:12-84 with the reference run's own projection directions (equal counts: sorted RMS;
    unequal: linear-interpolation quantiles)."""
    from mcmc_b200.benchmarks.metrics import sliced_wasserstein_distance
    got = sliced_wasserstein_distance(BLOB[f"sw2|{tag}|a"], BLOB[f"sw2|{tag}|b"], n_projections=30, projection_batch_size=16,
                                      directions=BLOB[f"sw2|{tag}|directions"])
    assert abs(got - float(BLOB[f"sw2|{tag}|value"].ravel()[0])) <= 1e-12


def test_detect_divergences_matches_reference_source(gpu_lib):
    from mcmc_b200.diagnostics import detect_divergences
    got = detect_divergences(torch.from_numpy(BLOB["divergences|dh"]).cuda())
    assert np.array_equal(got.cpu().numpy().astype(np.int8), BLOB["divergences|flags"])
