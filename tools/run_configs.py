"""Runs BASELINE.json's configs (other than the bench.py headline) through the public API and prints one
JSON line per measurement; `--out` collects them in a file for profiles/.

  1  README quickstart: GRAHMC tanh, 10-D standard normal (-0.5*sum x^2), 4 chains, 1000 samples, 500 burn-in
  2  HMC rosenbrock 65,536 chains: min-ESS/sec per L (the second half of the metric; bench.py reports grad-evals/s)
  3  GRAHMC x five schedules on the 20-D funnel: full windowed warmup (dual averaging + Welford) per GPU shard
  4  RWMH on the 20-D standard normal, 4M chains in total, 128 stored iterations (HBM-write path)
  5  NUTS, 65,536 chains, three targets, then R-hat / ESS / SW2

Under torchrun every rank runs its chain shard (global chain ids via chain_offset) and the cross-chain
statistics go through mcmc_b200/dist.py (NCCL).
"""
import argparse
import json
import math
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from mcmc_b200 import diagnostics as DG, dist  # noqa: E402
from mcmc_b200.benchmarks import targets as T  # noqa: E402
from mcmc_b200.benchmarks.metrics import compute_sliced_w2  # noqa: E402
from mcmc_b200.random import normal, split  # noqa: E402
from mcmc_b200.samplers import get_friction_schedule, hmc_run, nuts_run, rahmc_run, rwMH_run  # noqa: E402
from mcmc_b200.tuning import run_adaptive_warmup  # noqa: E402

D = 20
OUT = []


def emit(**kw):
    kw["n_gpus"] = dist.world_size()
    if dist.rank() == 0:
        print(json.dumps(kw), flush=True)
        OUT.append(kw)


class Timer:
    """Device time of a region (CUDA events on the current stream), max over ranks."""

    def __enter__(self):
        self.e0, self.e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        self.e0.record()
        return self

    def __exit__(self, *a):
        self.e1.record()
        torch.cuda.synchronize()
        t = torch.tensor([self.e0.elapsed_time(self.e1)], dtype=torch.float64, device="cuda")
        dist.all_reduce_max_(t)
        self.ms = float(t.item())


def shard(total):
    off, n = dist.chain_shard(total)
    return off, n


def config1():
    tgt = T.standard_normal(10, with_norm_const=False)          # README.md:51-52
    q0 = 2.0 * normal(0, (4, 10))
    args = dict(step_size=0.05, num_steps=16, gamma=1.0, steepness=0.5, num_samples=1000, burn_in=500,
                friction_schedule=get_friction_schedule("tanh"))
    rahmc_run(1, tgt.log_prob_fn, q0, **args)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    smp, lps, acc, st = rahmc_run(1, tgt.log_prob_fn, q0, **args)
    host = smp.cpu()
    dt = time.perf_counter() - t0
    emit(config=1, what="README quickstart GRAHMC tanh 10-D, 4 chains, 1000 samples + 500 burn-in, host arrays in/out",
         wall_ms=dt * 1e3, grad_evals=4 * 1500 * 16, mean_accept=float(acc.mean()), sample_mean_abs_max=float(host.mean((0, 1)).abs().max()))


def config2(chains_total):
    off, C = shard(chains_total)
    tgt = T.rosenbrock(D, scale=0.1)
    q0 = torch.from_numpy(1.0 + 0.5 * normal(100 + dist.rank(), (C, D))).cuda()
    w = torch.linspace(0.5, 2.0, D, dtype=torch.float64, device="cuda")
    for L in (16, 32, 64, 128):
        # same shapes as the timed call, so the 2.7 GB sample buffer comes from torch's cache, not cudaMalloc
        hmc_run(5, tgt.log_prob_fn, q0, 0.02, L, 256, burn_in=64, inv_mass_matrix=w, _chain_offset=off)
        with Timer() as tm:
            smp, lps, acc, st = hmc_run(7, tgt.log_prob_fn, q0, 0.02, L, 256, burn_in=64, inv_mass_matrix=w,
                                        _chain_offset=off)
        d = DG.compute_diagnostics(smp, distributed=True)
        emit(config=2, what="HMC 20-D rosenbrock(scale 0.1), eps 0.02, diag M^-1 linspace(0.5,2), 256 samples after 64 burn-in",
             L=L, chains=chains_total, ms=tm.ms, grad_evals_per_s=chains_total * 320 * L / (tm.ms * 1e-3),
             min_ess_bulk=d["ess_bulk_min"], min_ess_per_s=d["ess_bulk_min"] / (tm.ms * 1e-3), rhat_max=d["rhat_max"],
             mean_accept=float(dist.global_mean(acc.double().sum().reshape(1), C).item()))
        del smp, lps


def config3(chains_total, schedules, friction):
    off, C = shard(chains_total)
    tgt = T.neals_funnel(D)
    q0 = torch.from_numpy(tgt.init_sampler(split(3, 1 + dist.world_size())[dist.rank()], C)).cuda()
    # untimed pass with the same shapes: module load, torch's caching allocator, NCCL communicator warm-up
    run_adaptive_warmup("grahmc", tgt.log_prob_fn, None, q0, key=11, num_warmup=300, schedule_type=schedules[0],
                        num_steps=16, chain_offset=off, verbose=False, tune_friction=False)
    for sch in schedules:
        with Timer() as tm:
            eps, w, pos, info = run_adaptive_warmup("grahmc", tgt.log_prob_fn, None, q0, key=11, num_warmup=2500,
                                                    schedule_type=sch, num_steps=16, chain_offset=off, verbose=False,
                                                    tune_friction=friction)
        ge = torch.tensor([float(info["grad_evals_local"])], dtype=torch.float64, device="cuda")
        dist.all_reduce_sum_(ge)
        emit(config=3, what="GRAHMC funnel 20-D: run_adaptive_warmup, 2500 steps / 26 DA updates, device Welford"
                            + (" + phase-3 friction tuning" if friction else ""),
             schedule=sch, chains=chains_total, L=16, warmup_ms=tm.ms, warmup_grad_evals=float(ge.item()),
             warmup_grad_evals_per_s=float(ge.item()) / (tm.ms * 1e-3), step_size=eps,
             inv_mass_min=float(w.min()), inv_mass_max=float(w.max()), gamma=info["gamma"], steepness=info["steepness"])


def config4(chains_total, S):
    off, C = shard(chains_total)
    tgt = T.standard_normal(D)
    q0 = torch.from_numpy(0.1 * normal(200 + dist.rank(), (C, D))).cuda()
    scale = 2.38 / math.sqrt(D)
    rwMH_run(1, tgt.log_prob_fn, q0[:65536], 4, scale, burn_in=0, _chain_offset=off)
    for rep in range(2):
        with Timer() as tm:
            smp, lps, acc, st = rwMH_run(2 + rep, tgt.log_prob_fn, q0, S, scale, burn_in=16, _chain_offset=off)
        acc_mean = float(dist.global_mean(acc.double().sum().reshape(1), C).item())
        m = float(smp[-1].mean())
        del smp, lps          # the 90 GB block returns to torch's cache; the second (reported) pass reuses it
    bytes_written = chains_total * S * (D + 1) * 8
    emit(config=4, what="RWMH 20-D standard normal, scale 2.38/sqrt(20), 16 burn-in, samples (S,C,D)+log_probs stored",
         chains=chains_total, S=S, ms=tm.ms, stored_GB=bytes_written / 1e9, write_GBps=bytes_written / (tm.ms * 1e-3) / 1e9,
         chain_iters_per_s=chains_total * (S + 16) / (tm.ms * 1e-3), mean_accept=acc_mean, last_mean=m)


def config5(chains_total, S):
    off, C = shard(chains_total)
    cases = (("standard_normal", T.standard_normal(D), 0.5, lambda k: normal(k, (C, D))),
             ("rosenbrock", T.rosenbrock(D, scale=0.1), 0.02, lambda k: 1.0 + 0.5 * normal(k, (C, D))),
             ("neals_funnel", T.neals_funnel(D), 0.1, lambda k: T.neals_funnel(D).init_sampler(k, C)))
    for name, tgt, eps, init in cases:
        q0 = torch.from_numpy(init(300 + dist.rank())).cuda()
        nuts_run(1, tgt.log_prob_fn, q0, eps, S, burn_in=32, _chain_offset=off)   # same shapes: buffers come from torch's cache
        with Timer() as tm:
            smp, lps, acc, st, depths, maps = nuts_run(2, tgt.log_prob_fn, q0, eps, S, burn_in=32, _chain_offset=off)
        # grad-evals = sum 2^depth - 1 over transitions (run_benchmarks.py:724), counted over stored iterations here
        ge = ((2.0 ** depths.double()) - 1.0).sum().reshape(1)
        dsum = depths.double().sum().reshape(1)
        ge, dsum = dist.fused_sum([ge, dsum])
        with Timer() as td_:
            d = DG.compute_diagnostics(smp, distributed=True)
            z = DG.check_summary_statistics(d, tgt)
            sw2 = compute_sliced_w2(smp, name, D, distributed=True)
        emit(config=5, what="NUTS max_tree_depth 10, delta_max 1000, 32 burn-in", target=name, chains=chains_total, S=S,
             step_size=eps, ms=tm.ms, grad_evals_per_s=float(ge.item()) * (S + 32) / S / (tm.ms * 1e-3),
             mean_depth=float(dsum.item()) / (chains_total * S), max_depth=int(depths.max()),
             mean_accept_prob=float(dist.global_mean(maps.sum().reshape(1), C * S).item()),
             rhat_max=d["rhat_max"], ess_bulk_min=d["ess_bulk_min"], ess_tail_min=d["ess_tail_min"],
             min_ess_per_s=d["ess_bulk_min"] / (tm.ms * 1e-3), max_z=z.get("max_z"), z_threshold=z.get("threshold"),
             sliced_w2=sw2, diagnostics_ms=td_.ms)
        del smp, lps, depths, maps


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--configs", default="1,2,3,4,5")
    ap.add_argument("--out", default=None)
    ap.add_argument("--c3-chains", type=int, default=None, help="total chains of config 3 (default 32768 per GPU)")
    ap.add_argument("--c3-schedules", default="constant,tanh,sigmoid,linear,sine")
    ap.add_argument("--c3-friction", action="store_true", help="also run the phase-3 friction tuning")
    ap.add_argument("--c4-chains", type=int, default=4194304)
    ap.add_argument("--c4-samples", type=int, default=128)
    ap.add_argument("--c5-samples", type=int, default=256)
    a = ap.parse_args()
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        import torch.distributed as td
        td.init_process_group("nccl", device_id=torch.device("cuda", local))
        dist.warm_up()
    want = {int(x) for x in a.configs.split(",")}
    if 1 in want and dist.rank() == 0:
        config1()
    if 2 in want:
        config2(65536 * world)
    if 3 in want:
        config3(a.c3_chains or 32768 * world, a.c3_schedules.split(","), a.c3_friction)
    if 4 in want:
        config4(a.c4_chains, a.c4_samples)
    if 5 in want:
        config5(65536 * world, a.c5_samples)
    if a.out and dist.rank() == 0:
        with open(a.out, "w") as f:
            for r in OUT:
                f.write(json.dumps(r) + "\n")
    if world > 1:
        import torch.distributed as td
        td.destroy_process_group()


if __name__ == "__main__":
    main()
