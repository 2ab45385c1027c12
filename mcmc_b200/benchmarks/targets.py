"""Target distributions — host-side mirror of the reference's benchmarks/targets.py.

Same factories, same `TargetDistribution` fields (targets.py:23-31).  The
`log_prob_fn` of each target is a *tagged* callable: the samplers read its
`target_id` / `target_params` and run the analytic log-density + gradient that
is fused into the CUDA kernels (mcmc_b200/csrc/targets.cuh); calling it directly
evaluates the same device code on a batch of points.  An untagged Python
callable cannot be traced into a kernel and is rejected (`NotImplementedError`)
— there is no CPU fallback.
"""
from __future__ import annotations

import math
from typing import Callable, NamedTuple, Optional

import numpy as np

from .. import _abi


class TargetDistribution(NamedTuple):
    """Container for a target (reference: benchmarks/targets.py:23-31)."""
    log_prob_fn: Callable
    dim: int
    true_mean: np.ndarray
    true_cov: Optional[np.ndarray]
    name: str
    description: str
    init_sampler: Optional[Callable] = None  # (key, n_chains) -> positions


class TaggedLogProb:
    """log_prob_fn stand-in carrying the kernel tag (target id + parameters)."""

    def __init__(self, target_name: str, dim: int, params, cov_inv=None, target_id=None):
        self.target_name = target_name
        # target_id is given for user targets registered at run time (mcmc_b200/custom.py)
        self.target_id = _abi.TARGET_IDS[target_name] if target_id is None else int(target_id)
        self.dim = int(dim)
        self.target_params = [float(x) for x in params]
        self.cov_inv = None if cov_inv is None else np.ascontiguousarray(cov_inv, dtype=np.float64)

    def __call__(self, x):
        from ..runtime import target_eval  # device evaluation (needs torch + a GPU)
        return target_eval(self, x, want_grad=False)[0]

    def value_and_grad(self, x):
        from ..runtime import target_eval
        return target_eval(self, x, want_grad=True)

    def __repr__(self):
        return f"TaggedLogProb({self.target_name}, dim={self.dim}, params={self.target_params})"


def resolve_log_prob(log_prob_fn) -> TaggedLogProb:
    """Map a reference-style `log_prob_fn` argument to its kernel tag."""
    if isinstance(log_prob_fn, TaggedLogProb):
        return log_prob_fn
    if isinstance(log_prob_fn, TargetDistribution):
        return resolve_log_prob(log_prob_fn.log_prob_fn)
    tag = getattr(log_prob_fn, "mcmc_b200_tag", None)
    if isinstance(tag, TaggedLogProb):
        return tag
    raise NotImplementedError(
        "log_prob_fn must come from mcmc_b200.benchmarks.targets (its analytic log-density and gradient are "
        "fused into the CUDA kernels) or from mcmc_b200.custom.compile_target (value + gradient supplied as CUDA "
        "source); arbitrary Python callables cannot be compiled into the sm_100a path and there is no CPU fallback")


def _normal_init(std):
    def init_sampler(key, n_chains, dim):
        from ..random import normal
        return normal(key, (n_chains, dim)) * std
    return init_sampler


def standard_normal(dim: int = 10, with_norm_const: bool = True) -> TargetDistribution:
    """N(0, I) (targets.py:38-61).  with_norm_const=False gives README.md:51's `-0.5*sum(x**2)`."""
    return TargetDistribution(
        log_prob_fn=TaggedLogProb("standard_normal", dim, [1.0 if with_norm_const else 0.0]),
        dim=dim, true_mean=np.zeros(dim), true_cov=np.eye(dim),
        name=f"StandardNormal{dim}D",
        description=f"{dim}D standard normal N(0, I) - tests basic correctness")


def correlated_gaussian(dim: int = 10, correlation: float = 0.9) -> TargetDistribution:
    """Compound-symmetry Gaussian (targets.py:64-107)."""
    cov = (1.0 - correlation) * np.eye(dim) + correlation * np.ones((dim, dim))
    return TargetDistribution(
        log_prob_fn=TaggedLogProb("correlated_gaussian", dim, [correlation]),
        dim=dim, true_mean=np.zeros(dim), true_cov=cov,
        name=f"CorrelatedGaussian{dim}D_rho{correlation}",
        description=f"{dim}D Gaussian with correlation rho={correlation} - tests handling of correlation")


def ill_conditioned_gaussian(dim: int = 10, condition_number: float = 100.0) -> TargetDistribution:
    """Diagonal Gaussian, eigenvalues linspace(1, kappa, dim) (targets.py:110-146)."""
    eig = np.linspace(1.0, condition_number, dim)
    return TargetDistribution(
        log_prob_fn=TaggedLogProb("ill_conditioned_gaussian", dim, [condition_number]),
        dim=dim, true_mean=np.zeros(dim), true_cov=np.diag(eig),
        name=f"IllConditioned{dim}D_kappa{int(condition_number)}",
        description=f"{dim}D Gaussian with kappa={condition_number} - tests ill-conditioning")


def neals_funnel(dim: int = 10) -> TargetDistribution:
    """Neal's funnel (targets.py:149-213)."""
    def init_sampler(key, n_chains):
        from ..random import normal, split
        k1, k2 = split(key)
        x0 = normal(k1, (n_chains, 1)) * 3.0
        xr = normal(k2, (n_chains, dim - 1))
        return np.concatenate([x0, xr], axis=1)

    diag = np.concatenate([[9.0], np.ones(dim - 1) * math.exp(4.5)])
    return TargetDistribution(
        log_prob_fn=TaggedLogProb("neals_funnel", dim, []),
        dim=dim, true_mean=np.zeros(dim), true_cov=np.diag(diag),
        name=f"NealsFunnel{dim}D",
        description=f"{dim}D Neal's funnel - tests varying curvature and scale",
        init_sampler=init_sampler)


def log_gamma(dim: int = 10, shape: float = 2.0, rate: float = 1.0) -> TargetDistribution:
    """Independent Gamma(shape, rate) (targets.py:216-270)."""
    lognorm = math.lgamma(shape) + shape * math.log(rate)

    def init_sampler(key, n_chains):
        from ..random import gamma
        return gamma(key, shape, (n_chains, dim)) / rate

    return TargetDistribution(
        log_prob_fn=TaggedLogProb("log_gamma", dim, [shape, rate, lognorm]),
        dim=dim, true_mean=np.ones(dim) * (shape / rate), true_cov=np.eye(dim) * (shape / rate ** 2),
        name=f"LogGamma{dim}D_shape{shape}_rate{rate}",
        description=f"{dim}D independent Gamma - tests heavy tails and asymmetry",
        init_sampler=init_sampler)


def student_t(dim: int = 10, df: float = 3.0) -> TargetDistribution:
    """Independent Student-t(df) (targets.py:273-334)."""
    lognorm = math.lgamma((df + 1.0) / 2.0) - math.lgamma(df / 2.0) - 0.5 * math.log(df * math.pi)

    def init_sampler(key, n_chains):
        from ..random import normal
        return normal(key, (n_chains, dim)) * 2.0

    true_cov = np.eye(dim) * (df / (df - 2.0)) if df > 2 else None
    return TargetDistribution(
        log_prob_fn=TaggedLogProb("student_t", dim, [df, lognorm]),
        dim=dim, true_mean=np.zeros(dim), true_cov=true_cov,
        name=f"StudentT{dim}D_df{df}",
        description=f"{dim}D independent Student-t(df={df}) - tests heavy tails and non-Gaussian geometry",
        init_sampler=init_sampler)


def rosenbrock(dim: int = 10, scale: float = 0.1) -> TargetDistribution:
    """Rosenbrock density (targets.py:337-408)."""
    def init_sampler(key, n_chains):
        from ..random import normal
        return np.ones((n_chains, dim)) + normal(key, (n_chains, dim)) * 0.5

    return TargetDistribution(
        log_prob_fn=TaggedLogProb("rosenbrock", dim, [scale]),
        dim=dim, true_mean=np.ones(dim), true_cov=None,
        name=f"Rosenbrock{dim}D_scale{scale}",
        description=f"{dim}D Rosenbrock(scale={scale}) - tests curved valleys and non-linear geometry",
        init_sampler=init_sampler)


def gaussian_mixture(dim: int = 10, n_modes: int = 2, separation: float = 5.0) -> TargetDistribution:
    """Two-mode mixture in x[0] (targets.py:411-482)."""
    if n_modes != 2:
        raise NotImplementedError("Only 2-mode mixture currently supported")

    def init_sampler(key, n_chains):
        from ..random import normal, split
        key1, key2 = split(key)
        n_half = n_chains // 2
        # the reference draws both halves from key1 (targets.py:461-462); kept
        x0_1 = normal(key1, (n_half,)) - separation / 2.0
        x0_2 = normal(key1, (n_chains - n_half,)) + separation / 2.0
        x0 = np.concatenate([x0_1, x0_2])[:, None]
        xr = normal(key2, (n_chains, dim - 1))
        return np.concatenate([x0, xr], axis=1)

    diag = np.concatenate([[1.0 + (separation / 2.0) ** 2], np.ones(dim - 1)])
    return TargetDistribution(
        log_prob_fn=TaggedLogProb("gaussian_mixture", dim, [separation]),
        dim=dim, true_mean=np.zeros(dim), true_cov=np.diag(diag),
        name=f"GaussianMixture{dim}D_modes{n_modes}_sep{separation}",
        description=f"{dim}D Gaussian mixture (x[0] bimodal) - tests mode-switching",
        init_sampler=init_sampler)


def dense_gaussian(cov) -> TargetDistribution:
    """Zero-mean Gaussian with a general dense covariance: the `x @ cov_inv @ x` form of
    targets.py:94 for an arbitrary matrix (not in the reference's registry)."""
    cov = np.asarray(cov, dtype=np.float64)
    dim = cov.shape[0]
    cov_inv = np.linalg.inv(cov)
    _, logdet = np.linalg.slogdet(cov)
    return TargetDistribution(
        log_prob_fn=TaggedLogProb("dense_gaussian", dim, [float(logdet)], cov_inv=cov_inv),
        dim=dim, true_mean=np.zeros(dim), true_cov=cov,
        name=f"DenseGaussian{dim}D", description=f"{dim}D Gaussian with a dense covariance")


_FACTORIES = {
    "standard_normal": standard_normal,
    "correlated_gaussian": correlated_gaussian,
    "ill_conditioned_gaussian": ill_conditioned_gaussian,
    "student_t": student_t,
    "log_gamma": log_gamma,
    "rosenbrock": rosenbrock,
    "neals_funnel": neals_funnel,
    "gaussian_mixture": gaussian_mixture,
}


def get_target(name: str, dim: int = 10, **kwargs) -> TargetDistribution:
    """targets.py:489-530.  The five `rahmc_paper_targets` names of the reference's registry
    import a module that is absent from the reference tree and cannot be specified."""
    if name not in _FACTORIES:
        raise ValueError(f"Unknown target '{name}'. Available: {list(_FACTORIES.keys())}")
    return _FACTORIES[name](dim=dim, **kwargs)


def list_targets():
    return [f(10) for f in _FACTORIES.values()]


# ---------------------------------------------------------------------------------------------
# ground-truth samplers (targets.py:561-701) — host-side draws from the Philox stream

def get_reference_sampler(target_name: str, dim: int = 10, **kwargs):
    """(key, n) -> (n, dim) ground-truth draws, or None (targets.py:561-686).  Rosenbrock: the reference
    loads benchmarks/reference_samples/rosenbrock_{20,50}d.npy, which are absent from its tree
     -> None, as the reference itself returns when the file is missing."""
    from ..random import bernoulli, gamma, normal, split
    if target_name == "standard_normal":
        return lambda key, n: normal(key, (n, dim))
    if target_name == "correlated_gaussian":
        rho = kwargs.get("correlation", 0.9)
        Lc = np.linalg.cholesky((1.0 - rho) * np.eye(dim) + rho * np.ones((dim, dim)))
        return lambda key, n: normal(key, (n, dim)) @ Lc.T
    if target_name == "ill_conditioned_gaussian":
        scales = np.sqrt(np.linspace(1.0, kwargs.get("condition_number", 100.0), dim))
        return lambda key, n: normal(key, (n, dim)) * scales
    if target_name == "student_t":
        df = kwargs.get("df", 3.0)

        def sampler(key, n):
            k1, k2 = split(key)
            z = normal(k1, (n, dim))
            chi2 = gamma(k2, df / 2.0, (n, 1)) * 2.0
            return z / np.sqrt(chi2 / df)
        return sampler
    if target_name == "log_gamma":
        shape, rate = kwargs.get("shape", 2.0), kwargs.get("rate", 1.0)
        return lambda key, n: gamma(key, shape, (n, dim)) / rate
    if target_name == "neals_funnel":
        def sampler(key, n):
            k1, k2 = split(key)
            v = normal(k1, (n,)) * 3.0
            rest = normal(k2, (n, dim - 1)) * np.exp(v / 2.0)[:, None]
            return np.concatenate([v[:, None], rest], axis=1)
        return sampler
    if target_name == "gaussian_mixture":
        if kwargs.get("n_modes", 2) != 2:
            return None
        sep = kwargs.get("separation", 5.0)

        def sampler(key, n):
            k1, k2, k3 = split(key, 3)
            comp = bernoulli(k1, 0.5, (n,))
            x0 = normal(k2, (n,)) + np.where(comp, sep / 2.0, -sep / 2.0)
            return np.concatenate([x0[:, None], normal(k3, (n, dim - 1))], axis=1)
        return sampler
    return None


def has_reference_sampler(target_name: str) -> bool:
    """targets.py:689-701."""
    return target_name in ("standard_normal", "correlated_gaussian", "ill_conditioned_gaussian", "student_t",
                           "log_gamma", "neals_funnel", "gaussian_mixture", "rosenbrock")
