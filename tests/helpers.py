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


# ---------------------------------------------------------------------------------------------
# CPU stand-ins for the device calls of the tuning code, built on the oracle.  They let the
# host-side logic (dual averaging, windowed Welford warmup, chain sharding + gloo reductions)
# run in the CPU test suite; the product modules themselves never import the oracle.

def oracle_sample_batch(sampler, log_prob_fn, position, key, num_samples, *, step_size, num_steps=0,
                        inv_mass_matrix=None, gamma=0.0, steepness=1.0, schedule_id=0, max_tree_depth=10,
                        welford=None, welford_count=0.0, store_samples=False, track_proposals=False, chain_offset=0):
    import torch
    from mcmc_b200.benchmarks.targets import resolve_log_prob
    from mcmc_b200.random import as_key
    from oracle import oracle as O
    tag = resolve_log_prob(log_prob_fn)
    q0 = np.ascontiguousarray(position.numpy() if hasattr(position, "numpy") else position, dtype=np.float64)
    Cn, D = q0.shape
    flags = (_abi.FLAG_ACCUM_WELFORD if welford is not None else 0)
    p = O.make_params(tag.target_id, tag.target_params, Cn, D, num_samples, 0, num_steps, step_size, gamma=gamma,
                      steepness=steepness, schedule_id=schedule_id, rng_mode=_abi.RNG_PHILOX, seed=tuple(as_key(key)),
                      chain_offset=chain_offset, flags=flags, max_tree_depth=max_tree_depth)
    w = None if welford is None else (welford[0].numpy(), welford[1].numpy())
    im = None if inv_mass_matrix is None else np.asarray(inv_mass_matrix, dtype=np.float64)
    r = O.run(sampler, p, q0, inv_mass=im, cov_inv=tag.cov_inv, welford=w, welford_count=welford_count)
    if welford is not None:   # the kernels update the caller's arrays in place
        welford[0].copy_(torch.from_numpy(r["welford_mean"]))
        welford[1].copy_(torch.from_numpy(r["welford_m2"]))
    out = {k: torch.from_numpy(np.ascontiguousarray(v.astype(np.int64) if v.dtype == np.uint64 else v))
           for k, v in r.items()}
    if sampler == _abi.SAMPLER_NUTS:
        out["accept_sum"] = out["mean_accept_probs"].sum()
    else:
        out["accept_sum"] = out["counters"][0].to(torch.float64)
    out["accept_terms"] = Cn * int(num_samples)
    return out


def oracle_welford_sum_var(m2, count):
    import torch
    from oracle import oracle as O
    return torch.from_numpy(O.welford_finalize(m2.numpy(), count))


def patch_tuning_to_oracle(monkeypatch=None):
    """Route the tuning modules' device calls to the oracle stand-ins (tests only)."""
    import torch
    import importlib
    adaptation = importlib.import_module("mcmc_b200.tuning.adaptation")
    dual_averaging = importlib.import_module("mcmc_b200.tuning.dual_averaging")
    seq = importlib.import_module("mcmc_b200.tuning.sequential_tune_grahmc")
    as_cpu = lambda x: torch.as_tensor(np.asarray(x), dtype=torch.float64).contiguous() if not isinstance(
        x, torch.Tensor) else x.to(torch.float64).contiguous()
    pairs = [(adaptation, "sample_batch", oracle_sample_batch), (adaptation, "welford_sum_var", oracle_welford_sum_var),
             (adaptation, "_as_state_tensor", as_cpu), (dual_averaging, "sample_batch", oracle_sample_batch),
             (seq, "sample_batch", oracle_sample_batch)]
    for mod, name, fn in pairs:
        if monkeypatch is not None:
            monkeypatch.setattr(mod, name, fn)
        else:
            setattr(mod, name, fn)
