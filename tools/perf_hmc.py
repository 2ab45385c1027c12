"""Kernel-level timing sweep of the HMC/GRAHMC/RWMH/NUTS kernels (CUDA events on the launching stream)."""
import argparse
import json
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from mcmc_b200 import _abi  # noqa: E402
from mcmc_b200.benchmarks import targets as T  # noqa: E402
from mcmc_b200.runtime import run_sampler  # noqa: E402


def timed(fn, reps=3):
    fn()
    torch.cuda.synchronize()
    best = 1e30
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        out = fn()
        e1.record()
        torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    return best, out


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--chains", type=int, default=65536)
    ap.add_argument("--samples", type=int, default=64)
    ap.add_argument("--target", default="rosenbrock")
    ap.add_argument("--Ls", default="16,32,64,128")
    ap.add_argument("--sampler", default="hmc")
    ap.add_argument("--eps", type=float, default=0.01)
    ap.add_argument("--no-store", action="store_true")
    a = ap.parse_args()
    D = 20
    tgt = T.get_target(a.target, D)
    tag = tgt.log_prob_fn
    rng = np.random.default_rng(0)
    q0 = torch.from_numpy(1.0 + 0.3 * rng.standard_normal((a.chains, D))).cuda()
    for L in [int(x) for x in a.Ls.split(",")]:
        smp = {"hmc": _abi.SAMPLER_HMC, "grahmc": _abi.SAMPLER_GRAHMC, "rwmh": _abi.SAMPLER_RWMH,
               "nuts": _abi.SAMPLER_NUTS}[a.sampler]
        fn = lambda: run_sampler(smp, tag, q0, key=1, step_size=a.eps, num_steps=L, num_samples=a.samples,
                                 burn_in=0, gamma=1.0, steepness=0.5, schedule_id=1,
                                 store_samples=not a.no_store)
        ms, out = timed(fn)
        cnt = out["counters"].cpu().numpy()
        ge = float(cnt[1]) if a.sampler != "rwmh" else a.chains * a.samples
        print(json.dumps(dict(sampler=a.sampler, target=a.target, L=L, chains=a.chains, samples=a.samples, ms=ms,
                              grad_evals_per_s=ge / (ms * 1e-3), accept=float(out["accept_rate"].mean()),
                              store=not a.no_store)))
