"""Serial latency of one NUTS chain: with few chains nothing can overlap, so time / (leapfrog steps per chain) is the
per-step latency a chain on the critical path sees.  python tools/nuts_latency_probe.py  (MCMC_NUTS_THREAD=1 selects the
thread-per-chain kernel in the fast build)."""
import os, sys, time
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from mcmc_b200.benchmarks import targets as T
from mcmc_b200.random import normal
from mcmc_b200.samplers import nuts_run
D = 20
for name, tgt, eps in (("standard_normal", T.standard_normal(D), 0.05), ("rosenbrock", T.rosenbrock(D, scale=0.1), 0.005),
                       ("neals_funnel", T.neals_funnel(D), 0.02)):
    for C in (1, 8, 32, 1024):
        q0 = torch.from_numpy(normal(1, (C, D)) * 0.3 + (1.0 if name == "rosenbrock" else 0.0)).cuda()
        best = 1e9
        for rep in range(3):
            torch.cuda.synchronize(); t0 = time.perf_counter()
            out = nuts_run(2, tgt.log_prob_fn, q0, eps, 24, burn_in=0)
            torch.cuda.synchronize(); best = min(best, time.perf_counter() - t0)
        dep = out[4].cpu().numpy().astype(np.int64)
        steps = (2 ** dep - 1).sum(0)                      # per chain
        print(f"{name:16s} C={C:5d} mean depth {dep.mean():.2f} max-chain steps {steps.max():7d} time {best*1e3:8.3f} ms "
              f"-> {best*1e9/steps.max():7.1f} ns per step of the heaviest chain ({best*1e9/steps.max()*1.965:6.0f} cycles)", flush=True)
