"""Where does run_adaptive_warmup's time go?  Wraps sample_batch with CUDA events (device time of the sampler
kernel) and host timers, then prints per-phase sums: device time in the sampler, host time inside sample_batch
(launch path), and everything else (syncs, dual averaging, NCCL, Welford finalize)."""
import argparse
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from mcmc_b200.benchmarks import targets as T  # noqa: E402
from mcmc_b200.random import split  # noqa: E402
from mcmc_b200.tuning import adaptation as AD  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--sampler", default="grahmc")
ap.add_argument("--chains", type=int, default=32768)
ap.add_argument("--target", default="neals_funnel")
ap.add_argument("--schedule", default="tanh")
ap.add_argument("--reps", type=int, default=3)
a = ap.parse_args()

D = 20
tgt = T.get_target(a.target, D)
q0 = torch.from_numpy(tgt.init_sampler(split(3, 2)[0], a.chains)).cuda()
orig = AD.sample_batch
rec = []


def wrapped(*args, **kw):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter()
    e0.record()
    out = orig(*args, **kw)
    e1.record()
    rec.append((e0, e1, time.perf_counter() - t0, args[4]))
    return out


AD.sample_batch = wrapped
for rep in range(a.reps):
    rec.clear()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    eps, w, pos, info = AD.run_adaptive_warmup(a.sampler, tgt.log_prob_fn, None, q0, key=11, num_warmup=2500,
                                               schedule_type=a.schedule if a.sampler == "grahmc" else None,
                                               num_steps=16, verbose=False, tune_friction=False)
    torch.cuda.synchronize()
    wall = (time.perf_counter() - t0) * 1e3
    dev = sum(e0.elapsed_time(e1) for e0, e1, _, _ in rec)
    host = sum(h for _, _, h, _ in rec) * 1e3
    print(f"rep {rep}: wall {wall:.1f} ms | {len(rec)} batches, device time in sampler {dev:.1f} ms, "
          f"host time inside sample_batch {host:.1f} ms, rest {wall - max(dev, host):.1f} ms | eps {eps:.4f}")
    if rep == a.reps - 1:
        for e0, e1, h, ns in rec:
            print(f"   batch S={ns}: device {e0.elapsed_time(e1):.3f} ms, host {h * 1e3:.3f} ms")
