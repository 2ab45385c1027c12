"""Host-side cost of one hmc_run call (wall clock, no sync) next to its kernel time: shows whether the
device-resident loop of bench.py is launch-bound."""
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from mcmc_b200.benchmarks.targets import rosenbrock  # noqa: E402
from mcmc_b200.samplers import hmc_run  # noqa: E402

D, C, S = 20, 65536, 64
tgt = rosenbrock(D, scale=0.1)
q0 = torch.from_numpy(1.0 + 0.5 * np.random.default_rng(0).standard_normal((C, D))).cuda()
w = torch.linspace(0.5, 2.0, D, dtype=torch.float64, device="cuda")
for L in (16, 32, 64, 128):
    for _ in range(3):
        out = hmc_run(1, tgt.log_prob_fn, q0, 0.02, L, S, inv_mass_matrix=w)
    torch.cuda.synchronize()
    n = 20
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter()
    e0.record()
    for i in range(n):
        out = hmc_run(i, tgt.log_prob_fn, q0, 0.02, L, S, inv_mass_matrix=w)
    e1.record()
    host = (time.perf_counter() - t0) / n
    torch.cuda.synchronize()
    print(f"L={L}: host {host*1e3:.3f} ms/call, device {e0.elapsed_time(e1)/n:.3f} ms/call", flush=True)
