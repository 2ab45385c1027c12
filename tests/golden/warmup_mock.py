"""A deterministic stand-in for `*_run` inside run_adaptive_warmup, shared by the fixture generator (which plugs it into
the REFERENCE's tuning/adaptation.py) and by tests/test_reference_host_golden_cpu.py (which plugs it into
mcmc_b200.tuning.adaptation).  With the sampler replaced by the same function on both sides, the final step size and
the learned metric depend only on the warmup control flow: schedule, batch sizes, dual-averaging updates and resets,
per-chain Welford accumulation, cross-chain mean of variances, shrinkage and floor."""
import numpy as np


def mock_batch(call_idx, position, step_size, inv_mass, num_samples):
    """-> samples (S, C, D) and integer accept counts (C,).  The accept rate falls with the step size (so dual averaging
    has something to track); the draws' spread depends on the current metric (so one window feeds the next)."""
    position = np.asarray(position, dtype=np.float64)
    C, D = position.shape
    w = np.ones(D) if inv_mass is None else np.asarray(inv_mass, dtype=np.float64)
    rng = np.random.default_rng(1000 + call_idx)
    noise = rng.standard_normal((num_samples, C, D))
    sd = (0.5 + 0.25 * np.arange(D)) * (1.0 + 0.1 * np.sqrt(w))
    samples = 0.2 * position[None] + (np.arange(C)[None, :, None] * 0.05) + noise * sd
    p_acc = np.clip(np.exp(-4.0 * float(step_size)) + 0.03 * (np.arange(C) - C / 2) / C, 0.0, 1.0)
    counts = np.floor(p_acc * num_samples + rng.random(C)).astype(np.int64)
    return samples, np.minimum(counts, num_samples)
