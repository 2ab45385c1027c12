"""Shared inputs of the dense_gaussian tests (target id 8: log p = -1/2 (x^T P x + logdet Sigma + D log 2pi) with a
general dense P — the `x @ cov_inv @ x` form of the reference's correlated_gaussian, benchmarks/targets.py:90-97)."""
import numpy as np


def compound_symmetry_cov(D, rho):
    """The covariance the reference's correlated_gaussian builds (targets.py:79-81)."""
    return (1.0 - rho) * np.eye(D) + rho * np.ones((D, D))


def random_spd(D, seed, cond=30.0):
    """A rotated diagonal covariance with eigenvalues in [1/cond, 1]: dense, SPD, not symmetric-structured."""
    rng = np.random.default_rng(seed)
    q, _ = np.linalg.qr(rng.standard_normal((D, D)))
    lam = np.exp(np.linspace(0.0, -np.log(cond), D))
    return (q * lam) @ q.T


def dense_target(cov):
    from mcmc_b200.benchmarks import targets as T
    return T.dense_gaussian(cov)
