"""How much of a NUTS warp's time is tail?  For thread-per-chain with a fixed lane->chain map a warp runs until its
slowest lane finishes: efficiency = mean over warps of (mean lane steps / max lane steps), from the stored tree depths."""
import os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from mcmc_b200.benchmarks import targets as T
from mcmc_b200.random import normal
from mcmc_b200.samplers import nuts_run
D, C = 20, 65536
for name, tgt, eps, q0 in (("standard_normal", T.standard_normal(D), 0.5, normal(1, (C, D))),
                           ("rosenbrock", T.rosenbrock(D, scale=0.1), 0.02, 1.0 + 0.5 * normal(2, (C, D))),
                           ("neals_funnel", T.neals_funnel(D), 0.1, T.neals_funnel(D).init_sampler(3, C))):
    for S in (32, 288):
        out = nuts_run(2, tgt.log_prob_fn, torch.from_numpy(q0).cuda(), eps, S, burn_in=0)
        dep = out[4].cpu().numpy().astype(np.int64)            # (S, C)
        steps = (2 ** dep - 1) + dep + 1                       # leapfrog trips + ~bookkeeping-only trips
        lane = steps.sum(0).reshape(-1, 32)
        print(f"{name:16s} S={S:3d} mean depth {dep.mean():.2f}  steps/transition {(2 ** dep - 1).mean():.1f}  "
              f"lane-time mean/max per warp: {(lane.mean(1) / lane.max(1)).mean():.3f}")
