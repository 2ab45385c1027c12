"""dense_gaussian (target id 8) on the CPU side: both oracles against the reference's own source.  With P the
compound-symmetry precision the dense product IS the reference's correlated_gaussian as coded (`x @ cov_inv @ x`,
benchmarks/targets.py:94), so its value and autodiff gradient are pinned by the reference-source fixtures; a random
dense SPD matrix is checked between the two restatements and against torch autograd of the same expression."""
import numpy as np
import pytest
import torch

from dense_cases import compound_symmetry_cov, dense_target, random_spd
from helpers import relerr
from mcmc_b200 import _abi
from mcmc_b200.benchmarks import targets as T
from oracle import oracle_np as ONP
from test_reference_host_golden_cpu import BLOB, TARGET_KEYS, parse_target_key

CORR_KEYS = [k for k in TARGET_KEYS if "|correlated_gaussian|" in k]


@pytest.mark.parametrize("key", CORR_KEYS)
def test_oracles_dense_gaussian_equal_reference_correlated_gaussian(oracle, key):
    _, D, kw = parse_target_key(key)
    tag = T.resolve_log_prob(dense_target(compound_symmetry_cov(D, kw["correlation"])).log_prob_fn)
    assert tag.target_id == _abi.TARGET_IDS["dense_gaussian"]
    x = BLOB[key + "|x"]
    p = oracle.make_params(tag.target_id, tag.target_params, x.shape[0], D, 0)
    lp, g = oracle.target_eval(p, x, cov_inv=tag.cov_inv)
    assert relerr(lp, BLOB[key + "|lp"]) <= 1e-12
    assert relerr(g, BLOB[key + "|grad"]) <= 1e-11
    lp2, g2 = ONP.dense_gaussian(x, tag.cov_inv, tag.target_params[0])
    assert relerr(lp2, BLOB[key + "|lp"]) <= 1e-12 and relerr(g2, BLOB[key + "|grad"]) <= 1e-11


@pytest.mark.parametrize("D", [20, 10, 7, 33])
def test_oracle_dense_gaussian_random_spd_vs_autograd(oracle, D):
    cov = random_spd(D, seed=D)
    t = dense_target(cov)
    tag = T.resolve_log_prob(t.log_prob_fn)
    x = np.random.default_rng(1).standard_normal((9, D))
    p = oracle.make_params(tag.target_id, tag.target_params, 9, D, 0)
    lp, g = oracle.target_eval(p, x, cov_inv=tag.cov_inv)
    xt = torch.tensor(x, dtype=torch.float64, requires_grad=True)
    P = torch.tensor(tag.cov_inv)
    v = -0.5 * (torch.einsum("ni,ij,nj->n", xt, P, xt) + tag.target_params[0] + D * np.log(2 * np.pi))
    v.sum().backward()
    assert relerr(lp, v.detach().numpy()) <= 1e-12 and relerr(g, xt.grad.numpy()) <= 1e-11
    np.testing.assert_allclose(t.true_cov, cov)
    assert abs(tag.target_params[0] - np.linalg.slogdet(cov)[1]) < 1e-12
