"""Per-call event times of hmc_run when calls with different L alternate (bench.py's loop)."""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from mcmc_b200.benchmarks.targets import rosenbrock  # noqa: E402
from mcmc_b200.samplers import hmc_run  # noqa: E402

D, C, S = 20, 65536, 64
tgt = rosenbrock(D, scale=0.1)
q0 = torch.from_numpy(1.0 + 0.5 * np.random.default_rng(0).standard_normal((C, D))).cuda()
w = torch.linspace(0.5, 2.0, D, dtype=torch.float64, device="cuda")


def loop(order, steps=5, keep=True):
    for L in order:
        out = hmc_run(0, tgt.log_prob_fn, q0, 0.02, L, S, inv_mass_matrix=w)
    torch.cuda.synchronize()
    evs = []
    for i in range(steps):
        for L in order:
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            o = hmc_run(np.array([i, L], dtype=np.uint32), tgt.log_prob_fn, q0, 0.02, L, S, inv_mass_matrix=w)
            if keep:
                out = o
            e1.record()
            evs.append((L, e0, e1))
    torch.cuda.synchronize()
    acc = {}
    for L, e0, e1 in evs:
        acc.setdefault(L, []).append(e0.elapsed_time(e1))
    return {L: round(float(np.mean(v)), 3) for L, v in acc.items()}


for name, order in (("grid", (16, 32, 64, 128)), ("reversed", (128, 64, 32, 16)), ("16/32", (16, 32)),
                    ("32 only", (32,)), ("32,32,16", (32, 32, 16)), ("64,32", (64, 32))):
    print(name, loop(order), flush=True)
print("grid, outputs dropped at once", loop((16, 32, 64, 128), keep=False), flush=True)
