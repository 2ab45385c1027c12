"""Trajectory-length grid search on top of the CUDA samplers: the reference's
`run_single_benchmark_with_L` (run_benchmarks.py:468-935) and `run_trajectory_length_grid_search`
(run_benchmarks.py:198-465) with the same argument order, result keys and selection rule.

Per L: warmup (run_adaptive_warmup for hmc / nuts / grahmc, dual_averaging_tune_rwmh for rwmh) -> one *_run
(burn_in 0, from the warmup's final positions) -> device diagnostics -> Bonferroni z-test -> usable / quality gates
-> Sliced-W2.  Selection: argmax of ess_per_gradient among the quality_pass runs, else among the usable runs
(:402-413); with no usable run an explicit failure record built from the "least bad" run (lowest R-hat, then highest
ESS, :264-273).  Gradient counts follow :717-731 (HMC/GRAHMC num_samples*L*n_chains, NUTS sum(2^depth - 1), RWMH 0).

Differences from the reference, all stated in the result dict: `sample_time` is DEVICE time of the sampling launch (CUDA
events; the reference's wall clock includes jit compilation), `min_ess_per_sec` = ess_bulk_min / sample_time is added
(the second half of BASELINE.json's metric), and `n_chains` is this process's shard when `chain_offset` / a process
group is used (diagnostics are then over all ranks' chains).  Not built: `track_convergence` (:572-655, a plotting aid)
raises NotImplementedError when requested.
"""
from __future__ import annotations

import time
import traceback
from typing import Dict, List, Optional

import numpy as np
import torch

from .. import diagnostics as DG
from .. import dist
from ..random import normal, split
from .metrics import compute_sliced_w2
from .targets import TargetDistribution, get_reference_sampler

_GRAHMC = ("grahmc", "rahmc")
# keys copied from the least-bad run into a failure record / from every run into grid_search_info["all_results"]
_FAILURE_KEYS = ("total_samples", "n_gradients", "rhat_max", "rhat_mean", "ess_bulk_min", "ess_bulk_mean", "ess_tail_min",
                 "ess_tail_mean", "ess_per_sample", "ess_per_gradient", "divergence_rate", "total_divergences",
                 "accept_rate", "warmup_time", "sample_time", "total_time", "sliced_w2", "stats_pass", "z_score_max",
                 "z_score_threshold", "convergence_trace", "rhat_pass", "ess_pass", "ess_tail_pass", "is_inefficient",
                 "is_high_efficiency", "step_size", "gamma", "steepness", "avg_tree_depth", "mass_matrix_learned",
                 "mass_matrix_min", "mass_matrix_max", "mass_matrix_mean")
_ROW_KEYS = ("num_steps", "ess_per_gradient", "ess_bulk_min", "ess_tail_min", "rhat_max", "rhat_mean", "accept_rate",
             "step_size", "total_samples", "n_gradients", "warmup_time", "sample_time", "usable", "quality_pass",
             "divergence_rate", "error", "sliced_w2", "z_score_max", "gamma", "steepness", "min_ess_per_sec")
_ROW_DEFAULTS = {"ess_per_gradient": 0, "ess_bulk_min": 0, "ess_tail_min": 0, "rhat_max": float("inf"), "total_samples": 0,
                 "n_gradients": 0, "usable": False, "quality_pass": False}


def _say(verbose, *a):
    if verbose and dist.rank() == 0:
        print(*a)


def _device_timed(fn):
    """(result, seconds of device time) of one launch sequence on the current stream."""
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    out = fn()
    e1.record()
    e1.synchronize()
    return out, e0.elapsed_time(e1) * 1e-3


def run_single_benchmark_with_L(sampler: str, target: TargetDistribution, target_name: str, key, n_chains: int,
                                num_warmup: int, num_samples: int, schedule_type: str, num_steps: int,
                                learn_mass_matrix: bool = True, track_convergence: bool = False,
                                convergence_base: float = 1.5, *, chain_offset: int = 0, distributed: bool = False,
                                verbose: bool = True) -> Dict:
    """One warmup -> sample -> gates run at a fixed trajectory length (run_benchmarks.py:468-935)."""
    from ..samplers import get_friction_schedule, hmc_run, nuts_run, rahmc_run, rwMH_run
    from ..tuning import dual_averaging_tune_rwmh, run_adaptive_warmup
    if track_convergence:
        raise NotImplementedError("track_convergence (run_benchmarks.py:572-655) is a plotting aid and is not built")
    start = time.time()
    sched = schedule_type if sampler in _GRAHMC else None
    try:
        key, init_key = split(key)
        if target.init_sampler is not None:                                   # :501-505
            init_pos = np.asarray(target.init_sampler(init_key, n_chains))
        else:
            init_pos = normal(init_key, (n_chains, target.dim)) * 0.1
        init_pos = torch.as_tensor(init_pos, dtype=torch.float64).cuda()

        t_w = time.time()
        if sampler == "rwmh":                                                 # :517-526
            step_size, _ = dual_averaging_tune_rwmh(key, target.log_prob_fn, init_pos, target_accept=0.234, max_iter=1000,
                                                    _chain_offset=chain_offset)
            inv_mass, pos, info = None, init_pos, {"scale": step_size}
        elif sampler in ("hmc", "nuts") + _GRAHMC:                            # :528-555
            kw = {"num_steps": num_steps} if sampler != "nuts" else {"max_tree_depth": 15}
            step_size, inv_mass, pos, info = run_adaptive_warmup(
                "grahmc" if sampler in _GRAHMC else sampler, target.log_prob_fn, None, init_pos, key,
                num_warmup=num_warmup, target_accept=0.65, schedule_type=sched, learn_mass_matrix=learn_mass_matrix,
                chain_offset=chain_offset, verbose=False, **kw)
        else:
            raise ValueError(f"Unknown sampler: {sampler}")
        torch.cuda.synchronize()
        warmup_time = time.time() - t_w

        key, sample_key = split(key)
        depths = None
        common = dict(_chain_offset=chain_offset)
        if sampler == "rwmh":                                                 # :662-668
            run = lambda: rwMH_run(sample_key, target.log_prob_fn, pos, num_samples, step_size, burn_in=0, **common)
        elif sampler == "hmc":                                                # :670-677
            run = lambda: hmc_run(sample_key, target.log_prob_fn, pos, step_size, num_steps, num_samples, burn_in=0,
                                  inv_mass_matrix=inv_mass, **common)
        elif sampler == "nuts":                                               # :679-686 (production depth cap 10)
            run = lambda: nuts_run(sample_key, target.log_prob_fn, pos, step_size, num_samples, burn_in=0,
                                   inv_mass_matrix=inv_mass, max_tree_depth=10, **common)
        else:                                                                 # :688-701
            run = lambda: rahmc_run(sample_key, target.log_prob_fn, pos, step_size, num_steps, info.get("gamma", 1.0),
                                    info.get("steepness", 5.0), num_samples, burn_in=0, inv_mass_matrix=inv_mass,
                                    friction_schedule=get_friction_schedule(schedule_type), **common)
        out, sample_time = _device_timed(run)
        samples, accept_rate = out[0], out[2]
        if sampler == "nuts":
            depths = out[4]
        world = dist.world_size() if distributed else 1
        acc = float(dist.global_mean(accept_rate.double().sum().reshape(1), n_chains).item()) if world > 1 \
            else float(accept_rate.double().mean())

        # gradient count (:717-731)
        avg_tree_depth = None
        if sampler == "nuts":
            parts = [((2.0 ** depths.double()) - 1.0).sum().reshape(1), depths.double().sum().reshape(1)]
            ge, dsum = dist.fused_sum(parts) if world > 1 else parts
            n_gradients = int(ge.item())
            avg_tree_depth = float(dsum.item()) / (num_samples * n_chains * world)
        elif sampler == "rwmh":
            n_gradients = 0
        else:
            n_gradients = num_samples * num_steps * n_chains * world

        if sampler == "rwmh":                                                 # sampler metadata (:734-752)
            meta = {"scale": step_size}
        elif sampler == "hmc":
            meta = {"step_size": step_size, "num_steps": num_steps}
        elif sampler == "nuts":
            meta = {"step_size": step_size, "max_tree_depth": 10, "avg_tree_depth": avg_tree_depth}
        else:
            meta = {"step_size": step_size, "num_steps": num_steps, "gamma": info.get("gamma", 1.0),
                    "steepness": info.get("steepness", 5.0), "schedule": schedule_type}

        d = DG.compute_diagnostics(samples, distributed=world > 1)            # :755
        stats = DG.check_summary_statistics(d, target, significance=0.05)     # :758
        has_truth = target.true_mean is not None and target.true_cov is not None
        divergence_rate, total_divergences = 0.0, 0                           # placeholder in the reference too (:706-710)
        g = DG.quality_gates(d, num_samples, divergence_rate=divergence_rate, stats_pass=stats["pass"],
                             has_true_mean=has_truth)                         # :777-804
        total_time = time.time() - start                                      # before SW2, as :807
        sliced_w2 = None
        if get_reference_sampler(target_name, target.dim) is not None:        # :811-825
            key, w2_key = split(key)
            sliced_w2 = compute_sliced_w2(samples, target_name, target.dim, n_reference=50000, n_projections=500,
                                          key=w2_key, distributed=world > 1)
        ess_min = d["ess_bulk_min"]
        res = {
            "sampler": sampler, "target": target.name, "schedule": sched, "dim": target.dim,
            "num_steps": num_steps if sampler in ("hmc",) + _GRAHMC else None, "n_chains": n_chains,
            "num_warmup": num_warmup, "num_samples": num_samples, "total_samples": num_samples,
            "warmup_time": warmup_time, "sample_time": sample_time, "total_time": total_time, "accept_rate": acc,
            "rhat_max": d["rhat_max"], "rhat_mean": d["rhat_mean"], "ess_bulk_min": ess_min,
            "ess_bulk_mean": d["ess_bulk_mean"], "ess_tail_min": d["ess_tail_min"], "ess_tail_mean": d["ess_tail_mean"],
            "ess_per_sample": g["ess_per_sample"], "ess_per_gradient": ess_min / n_gradients if n_gradients > 0 else 0,
            "min_ess_per_sec": ess_min / sample_time if sample_time > 0 else 0.0,
            "divergence_rate": divergence_rate, "total_divergences": total_divergences, "n_gradients": n_gradients,
            "rhat_pass": d["rhat_max"] < 1.01, "ess_pass": ess_min >= DG.MIN_ESS_QUALITY,
            "ess_tail_pass": d["ess_tail_min"] >= DG.MIN_ESS_TAIL_QUALITY, "stats_pass": stats["pass"],
            "z_score_max": stats.get("max_z"), "z_score_threshold": stats.get("threshold"),
            "usable": g["usable"], "quality_pass": g["quality_pass"], "is_inefficient": g["is_inefficient"],
            "is_high_efficiency": g["is_high_efficiency"], "sliced_w2": sliced_w2, "convergence_trace": None,
        }
        res.update(meta)
        res["mass_matrix_learned"] = learn_mass_matrix
        if learn_mass_matrix and inv_mass is not None:
            res["mass_matrix_min"] = float(inv_mass.min())
            res["mass_matrix_max"] = float(inv_mass.max())
            res["mass_matrix_mean"] = float(inv_mass.mean())
        status = "[PASS]" if res["quality_pass"] else ("[USABLE]" if res["usable"] else "[FAIL]")
        _say(verbose, f"  L={num_steps}: {status} R-hat {res['rhat_max']:.4f} | ESS {ess_min:.0f} | tail ESS "
                      f"{res['ess_tail_min']:.0f} | accept {acc:.3f} | eps {step_size:.5f} | "
                      f"sample {sample_time * 1e3:.2f} ms | min-ESS/s {res['min_ess_per_sec']:.3g}")
        return res
    except Exception as e:                                                    # :915-935: the sweep goes on
        if verbose:
            traceback.print_exc()
        return {"sampler": sampler, "target": target.name, "schedule": sched, "dim": target.dim, "num_steps": num_steps,
                "total_samples": 0, "ess_bulk_min": 0.0, "n_gradients": 0, "divergence_rate": None, "error": str(e),
                "total_time": time.time() - start, "usable": False, "quality_pass": False}


def _rows(results: List[Dict]) -> List[Dict]:
    return [{k: r.get(k, _ROW_DEFAULTS.get(k)) for k in _ROW_KEYS} for r in results]


def run_trajectory_length_grid_search(sampler: str, target: TargetDistribution, target_name: str, key, n_chains: int,
                                      num_warmup: int, num_samples: int, schedule_type: str, num_steps_grid: List[int],
                                      learn_mass_matrix: bool = True, track_convergence: bool = False,
                                      convergence_base: float = 1.5, *, chain_offset: int = 0, distributed: bool = False,
                                      verbose: bool = True) -> Dict:
    """Best run over `num_steps_grid` by ESS per gradient (run_benchmarks.py:198-465); the returned dict carries
    `grid_search_info` (every L's row, the selected L and the tier it was selected from)."""
    _say(verbose, f"GRID SEARCH: Testing trajectory lengths {num_steps_grid}")
    results = []
    for L in num_steps_grid:
        key, subkey = split(key)
        r = run_single_benchmark_with_L(sampler, target, target_name, subkey, n_chains, num_warmup, num_samples,
                                        schedule_type, L, learn_mass_matrix, track_convergence, convergence_base,
                                        chain_offset=chain_offset, distributed=distributed, verbose=verbose)
        if r.get("error") is None:                                            # :250-259
            n_grad = r.get("n_gradients", r["total_samples"] * L)
            r["n_gradients"] = n_grad
            r["ess_per_gradient"] = r["ess_bulk_min"] / n_grad if n_grad > 0 else 0
        else:
            r["n_gradients"], r["ess_per_gradient"] = 0, 0
        results.append(r)

    usable = [r for r in results if r.get("usable", False)]
    if not usable:                                                            # :264-400
        def badness(r):
            if r.get("error"):
                return (float("inf"), 0)
            return (r.get("rhat_max", float("inf")), -r.get("ess_bulk_min", 0))
        worst_ok = min(results, key=badness)
        _say(verbose, "!!! GRID SEARCH FAILED: NO TRAJECTORY LENGTH PRODUCED USABLE RESULTS !!!")
        fail = {"sampler": sampler, "target": target.name, "schedule": schedule_type if sampler in _GRAHMC else None,
                "dim": target.dim, "n_chains": n_chains, "grid_search_failed": True, "usable": False,
                "quality_pass": False, "error": "No trajectory length produced usable samples",
                "num_samples": num_samples, "num_warmup": num_warmup, "num_steps": worst_ok.get("num_steps")}
        for k in _FAILURE_KEYS:
            fail[k] = worst_ok.get(k, 0 if k in ("total_samples", "n_gradients", "ess_bulk_min", "ess_per_gradient") else None)
        fail["grid_search_info"] = {"tested_L_values": list(num_steps_grid), "selected_L": None, "has_usable": False,
                                    "least_bad_L": worst_ok.get("num_steps"), "all_results": _rows(results)}
        return fail

    quality = [r for r in usable if r.get("quality_pass", False)]            # :402-413
    tier, pool = ("quality_pass", quality) if quality else ("usable_only", usable)
    best = max(pool, key=lambda r: r["ess_per_gradient"])
    _say(verbose, f">>> BEST: L={best['num_steps']} (ESS/grad={best['ess_per_gradient']:.6f}) [{tier.upper()}]")
    best["grid_search_info"] = {"tested_L_values": list(num_steps_grid), "selected_L": best["num_steps"],
                                "selection_tier": tier, "has_usable": True, "all_results": _rows(results)}
    return best
