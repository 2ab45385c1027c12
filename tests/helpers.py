"""Shared test cases: every target with its kernel tag, an init distribution and a step size
at which the trajectories are well-conditioned (so 1e-10 parity is meaningful)."""
import math

import numpy as np

from mcmc_b200 import _abi

LGN = math.lgamma(2.0) + 2.0 * math.log(1.0)
STN = math.lgamma(2.0) - math.lgamma(1.5) - 0.5 * math.log(3.0 * math.pi)

# name -> (params, step_size, init(rng, C, D))
TARGET_CASES = {
    "standard_normal": ([1.0], 0.2, lambda r, C, D: r.standard_normal((C, D))),
    "correlated_gaussian": ([0.9], 0.05, lambda r, C, D: r.standard_normal((C, D))),
    "ill_conditioned_gaussian": ([100.0], 0.3, lambda r, C, D: r.standard_normal((C, D))),
    "neals_funnel": ([], 0.05, lambda r, C, D: r.standard_normal((C, D))),
    "log_gamma": ([2.0, 1.0, LGN], 0.05, lambda r, C, D: 0.5 + np.abs(r.standard_normal((C, D)))),
    "student_t": ([3.0, STN], 0.2, lambda r, C, D: 2.0 * r.standard_normal((C, D))),
    "rosenbrock": ([0.1], 0.005, lambda r, C, D: 1.0 + 0.05 * r.standard_normal((C, D))),
    "gaussian_mixture": ([5.0], 0.2, lambda r, C, D: r.standard_normal((C, D))),
}
POLYNOMIAL = ("standard_normal", "ill_conditioned_gaussian", "rosenbrock")  # no libm call on the path


def tid(name):
    return _abi.TARGET_IDS[name]


def relerr(a, b):
    a = np.asarray(a, dtype=np.float64)
    b = np.asarray(b, dtype=np.float64)
    if a.size == 0:
        return 0.0
    with np.errstate(invalid="ignore"):
        d = np.abs(a - b) / np.maximum(np.abs(b), 1.0)
    d = np.where((a == b) | (np.isnan(a) & np.isnan(b)), 0.0, d)
    return float(np.max(d))


class GpuBackend:
    """Adapter giving the CUDA path (through the C ABI's host entry) the oracle driver's surface."""

    def __init__(self, variant="fast"):
        self.variant = variant

    @staticmethod
    def make_params(*a, **k):
        from oracle.oracle import make_params  # same POD struct, same field meaning
        return make_params(*a, **k)

    def run(self, sampler, p, init_q, **kw):
        from mcmc_b200.hostrun import run_host
        kw.pop("store_samples", None)
        return run_host(sampler, p, init_q, variant=self.variant, **kw)
