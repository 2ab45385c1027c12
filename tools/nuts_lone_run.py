"""One NUTS run with a single warp of chains (for an ncu capture of the latency-bound regime)."""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from mcmc_b200.benchmarks import targets as T
from mcmc_b200.random import normal
from mcmc_b200.samplers import nuts_run
C = int(sys.argv[1]) if len(sys.argv) > 1 else 32
name = sys.argv[2] if len(sys.argv) > 2 else "neals_funnel"
tgt = getattr(T, name)(20)
q0 = torch.from_numpy(normal(1, (C, 20)) * 0.3).cuda()
if name == "rosenbrock":
    q0 = q0 * 0.2 + 1.0
for _ in range(2):
    out = nuts_run(2, tgt.log_prob_fn, q0, 0.02, 24, burn_in=0)
torch.cuda.synchronize()
print(float(out[4].double().mean()))
