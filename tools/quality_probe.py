"""Prints the gate statistics of a few sampler configurations (used to pick the settings of tests/test_quality_gpu.py)."""
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from mcmc_b200 import diagnostics as DG  # noqa: E402
from mcmc_b200.benchmarks import targets as T  # noqa: E402
from mcmc_b200.benchmarks.metrics import compute_sliced_w2  # noqa: E402
from mcmc_b200.random import normal  # noqa: E402
from mcmc_b200.samplers import hmc_run, nuts_run, rahmc_run, rwMH_run, get_friction_schedule  # noqa: E402

D = 20


def report(tag, samples, tgt, name, S, acc, **kw):
    t0 = time.time()
    d = DG.compute_diagnostics(samples)
    torch.cuda.synchronize()
    dt = time.time() - t0
    z = DG.check_summary_statistics(d, tgt)
    sw2 = compute_sliced_w2(samples, name, D, **kw)
    n = samples.shape[0] * samples.shape[1]
    print(f"{tag:34s} rhat {d['rhat_max']:.4f} essb/N {d['ess_bulk_min']/n:.3f} esst/N {d['ess_tail_min']/n:.3f} "
          f"maxz {z['max_z']:.2f} sw2 {sw2 if sw2 is None else round(sw2, 4)} acc {acc:.3f} diag {dt:.2f}s", flush=True)


C = 2048
sn = T.standard_normal(D)
q0 = torch.from_numpy(normal(9, (C, D))).cuda()
for eps, L in ((0.2, 8), (0.25, 6)):
    o = hmc_run(1, sn.log_prob_fn, q0, eps, L, 400, burn_in=100)
    report(f"hmc eps={eps} L={L}", o[0], sn, "standard_normal", 400, float(o[2].mean()))
for sch in ("constant", "tanh", "sigmoid", "linear", "sine"):
    for eps, L, g, S in ((0.2, 8, 0.5, 400), (0.2, 8, 0.1, 400), (0.25, 8, 0.2, 1000), (0.3, 6, 1.0, 1000)):
        st = 0.5 if sch == "tanh" else 2.0
        o = rahmc_run(2, sn.log_prob_fn, q0, eps, L, g, st, S, burn_in=100, friction_schedule=get_friction_schedule(sch))
        report(f"grahmc-{sch} e={eps} L={L} g={g} S={S}", o[0], sn, "standard_normal", S, float(o[2].mean()))
o = nuts_run(3, sn.log_prob_fn, q0, 0.5, 400, burn_in=100)
report("nuts eps=0.5", o[0], sn, "standard_normal", 400, float(o[2].mean()))
q1 = q0[:512]
for S in (4000, 12000):
    o = rwMH_run(4, sn.log_prob_fn, q1, S, 2.38 / np.sqrt(D), burn_in=1000)
    report(f"rwmh C=512 S={S}", o[0], sn, "standard_normal", S, float(o[2].mean()))
cg = T.correlated_gaussian(D, 0.9)
for eps, L in ((0.12, 24), (0.12, 48), (0.15, 40)):
    o = hmc_run(5, cg.log_prob_fn, q0, eps, L, 600, burn_in=300)
    report(f"corr hmc eps={eps} L={L}", o[0], cg, "correlated_gaussian", 600, float(o[2].mean()), correlation=0.9)
ic = T.ill_conditioned_gaussian(D, 100.0)
w = torch.linspace(1.0, 100.0, D, dtype=torch.float64, device="cuda")
for eps, L, ww in ((0.5, 24, None), (0.2, 8, w), (0.3, 5, w)):
    o = hmc_run(5, ic.log_prob_fn, q0 * (w.sqrt() if ww is not None else 1.0), eps, L, 600, burn_in=300, inv_mass_matrix=ww)
    report(f"illc hmc eps={eps} L={L} mass={'true' if ww is not None else 'I'}", o[0], ic, "ill_conditioned_gaussian", 600,
           float(o[2].mean()), condition_number=100.0)
