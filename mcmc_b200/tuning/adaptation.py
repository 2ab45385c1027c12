"""Windowed warmup: step size (dual averaging) + diagonal mass matrix (per-chain Welford),
the reference's `build_schedule` / `run_adaptive_warmup` (tuning/adaptation.py:17-348) over
the CUDA samplers.

What is different from the reference, and why results are the same:
  * the per-chain Welford states of an adaptation window are (C, D) arrays in HBM updated
    inside the sampler kernel with every stored iteration (MCMC_FLAG_ACCUM_WELFORD) — the
    same fold, in the same order, as the reference's double Python loop (:229-235);
  * samples of warmup batches are never stored (MCMC_FLAG_NO_SAMPLES), only the state moves on;
  * the two cross-chain statistics, mean acceptance (:218) and the mean over chains of the
    per-chain variances (:240-248), are reduced on the device and all-reduced over the chain
    shards of a multi-GPU run (mcmc_b200/dist.py); with one process they are the plain means.
"""
from __future__ import annotations

import math
import time
from typing import Any, Dict, Optional, Tuple

import torch

from .. import _abi, dist
from ..random import fold_in, split
from ..runtime import _device, to_device, welford_sum_var
from ..samplers.GRAHMC import _schedule_id, get_friction_schedule
from .dual_averaging import DeviceDualAveraging, da_init, da_reset, da_update, sample_batch



def _as_state_tensor(x) -> torch.Tensor:
    """init positions -> contiguous float64 CUDA tensor (no CPU fallback: raises without a device)."""
    return to_device(x, _device(x.device if getattr(x, "is_cuda", False) else None))


_SAMPLERS = {"hmc": _abi.SAMPLER_HMC, "nuts": _abi.SAMPLER_NUTS, "grahmc": _abi.SAMPLER_GRAHMC,
             "rahmc": _abi.SAMPLER_GRAHMC}


def build_schedule(num_steps: int = None, exploration_steps: int = 500, adaptation_windows: list = None,
                   cooldown_steps: int = 125) -> list:
    """[(start, end, phase)]: exploration, doubling adaptation windows, cooldown
    (tuning/adaptation.py:17-69; defaults total 500 + 1875 + 125 = 2500)."""
    if adaptation_windows is None:
        adaptation_windows = [25, 50, 100, 200, 500, 1000]
    schedule = []
    start = 0
    schedule.append((start, start + exploration_steps, "exploration"))
    start += exploration_steps
    for window_size in adaptation_windows:
        schedule.append((start, start + window_size, "adaptation"))
        start += window_size
    schedule.append((start, start + cooldown_steps, "cooldown"))
    start += cooldown_steps
    if num_steps is not None and start != num_steps and dist.rank() == 0:
        print(f"Warning: Computed warmup ({start}) doesn't match num_steps ({num_steps})")
        print(f"  Using computed warmup of {start} steps")
    return schedule


def _log(verbose, *a):
    if verbose and dist.rank() == 0:
        print(*a)


def run_adaptive_warmup(sampler: str, target_log_prob: Any, target_grad_log_prob: Any, initial_position,
                        key, num_warmup: int = 1000, target_accept: float = 0.65,
                        schedule_type: Optional[str] = None, update_freq: int = 100,
                        learn_mass_matrix: bool = True, **kwargs) -> Tuple[float, Optional[torch.Tensor], torch.Tensor, Dict]:
    """(step_size, inv_mass_matrix, final_position, info) — tuning/adaptation.py:72-348.

    `initial_position` is THIS rank's shard (n_chains_local, n_dim); pass `chain_offset=` (global index of
    its first chain) in a multi-GPU run so the Philox streams are those of the unsharded run.  Extra
    keyword arguments as in the reference: num_steps (default 20), max_tree_depth (default 10); plus
    `verbose` (default True) and, for GRAHMC, `tune_friction` (default True: phase 3).
    """
    if sampler not in _SAMPLERS:
        raise ValueError(f"Unknown sampler: {sampler}")
    sid = _SAMPLERS[sampler]
    verbose = kwargs.get("verbose", True)
    chain_offset = int(kwargs.get("chain_offset", 0))
    n_chains, n_dim = initial_position.shape
    start_time = time.time()

    initial_step = 0.5 / math.sqrt(n_dim)                       # adaptation.py:122
    grahmc = sid == _abi.SAMPLER_GRAHMC
    gamma = 1.0 if grahmc else None                              # :126-131
    steepness = (0.5 if schedule_type == "tanh" else 2.0) if grahmc else None
    sched_id = _schedule_id(get_friction_schedule(schedule_type or "constant")) if grahmc else 0

    position = _as_state_tensor(initial_position)
    dev = position.device
    # On a CUDA device the dual-averaging state lives there too (DeviceDualAveraging): the sampler reads epsilon from
    # device memory, the accept statistic goes kernel -> (all-reduce) -> da kernel, and nothing in the loop below waits
    # for the GPU.  The host state machine (same formulas) drives the CPU stand-ins of the test suite.
    on_device = position.is_cuda
    da_dev = DeviceDualAveraging(initial_step, dev) if on_device else None
    da_state = da_init(initial_step)
    inv_mass_matrix = None                                       # identity until the first window ends (:136)
    schedule = build_schedule(num_warmup)
    _log(verbose, f"Adaptation Schedule ({num_warmup} steps):")
    for s, e, t in schedule:
        _log(verbose, f"  [{s:4d} - {e:4d}] {t}")
    if not learn_mass_matrix:
        _log(verbose, "  [Mass matrix learning disabled - using identity]")

    num_steps = kwargs.get("num_steps", 20)
    max_tree_depth = kwargs.get("max_tree_depth", 10)
    wf_mean = wf_m2 = None
    wf_count = 0.0
    n_batches_total = 0
    grad_evals = 0
    grad_evals_dev = torch.zeros((), dtype=torch.int64, device=dev) if on_device else None
    for start_idx, end_idx, phase in schedule:
        window_len = end_idx - start_idx
        adapting = phase == "adaptation" and learn_mass_matrix
        if adapting:
            wf_count = 0.0                                       # fresh per-chain Welford states (:160-161)
            if wf_mean is not None:
                wf_mean.zero_()
                wf_m2.zero_()
        num_batches = max(1, window_len // update_freq)          # :164-165
        samples_per_batch = window_len // num_batches
        for _ in range(num_batches):
            key, subkey = split(key)
            current_step_size = 0.0 if on_device else math.exp(da_state.log_step)
            if adapting and wf_mean is None:
                wf_mean = torch.zeros((n_chains, n_dim), dtype=torch.float64, device=dev)
                wf_m2 = torch.zeros_like(wf_mean)
            out = sample_batch(sid, target_log_prob, position, subkey, samples_per_batch,
                               step_size=current_step_size, num_steps=num_steps, inv_mass_matrix=inv_mass_matrix,
                               gamma=gamma or 0.0, steepness=steepness or 1.0, schedule_id=sched_id,
                               max_tree_depth=max_tree_depth, welford=(wf_mean, wf_m2) if adapting else None,
                               welford_count=wf_count, chain_offset=chain_offset,
                               **({"step_size_dev": da_dev.step_dev} if on_device else {}))
            position = out["final_q"]
            if adapting:
                wf_count += samples_per_batch
            if on_device:
                # sum over ranks on the sampler's stream, then the one-thread da_update kernel (:217-219)
                # (the term count rides in the same payload: shards may differ in size by one chain)
                acc = torch.cat([out["accept_sum"].reshape(1),
                                 torch.full((1,), float(out["accept_terms"]), dtype=torch.float64, device=dev)])
                da_dev.update(dist.all_reduce_sum_(acc), 0.0, target_accept)
                grad_evals_dev += out["counters"][1]
            else:
                avg_accept = dist.mean_accept(out["accept_sum"], out["accept_terms"])   # :217-219
                da_state = da_update(da_state, avg_accept, target_accept)
                grad_evals += int(out["counters"][1].item())
            n_batches_total += 1
        if adapting:
            # mean over ALL chains of the per-chain variances, shrunk towards 1 (:238-265)
            local = welford_sum_var(wf_m2, wf_count)
            variance = dist.mean_chain_variance(local, n_chains)
            n_samples = wf_count
            shrinkage_weight = n_samples / (n_samples + 5.0)
            prior_weight = 5.0 / (n_samples + 5.0)
            variance = shrinkage_weight * variance + prior_weight * 1.0
            variance = torch.clamp(variance, min=1e-8)
            inv_mass_matrix = variance
            if verbose:      # reading the range synchronises: only when someone is watching
                _log(verbose, f"  Window finished. Updated Mass Matrix. Range: "
                              f"[{float(variance.min()):.4f}, {float(variance.max()):.4f}]")
                _log(verbose, f"    (n_samples_per_chain={n_samples:.0f}, n_chains={n_chains * dist.world_size()}, "
                              f"shrinkage={shrinkage_weight:.3f})")
            if on_device:
                da_dev.reset()                                   # :272-273
            else:
                da_state = da_reset(da_state)

    if on_device:
        da_state = da_dev.to_host()                              # the first (and only) wait for the GPU
        grad_evals = int(grad_evals_dev.item())
    final_step_size = float(math.exp(da_state.log_step_bar))
    _log(verbose, f"Warmup Complete. Final step_size: {final_step_size:.5f}")
    if inv_mass_matrix is None:
        inv_mass_matrix = torch.ones(n_dim, dtype=torch.float64, device=dev)

    # ---- phase 3: friction refinement on the learned metric (:279-318)
    tune_history = None
    if grahmc and kwargs.get("tune_friction", True):
        from .sequential_tune_grahmc import sequential_tune_grahmc
        _log(verbose, "\n[Phase 3] Tuning GRAHMC friction on learned mass matrix...")
        tuned_step, tuned_gamma, tuned_steepness, tune_history = sequential_tune_grahmc(
            key=fold_in(key, 999), log_prob_fn=target_log_prob, grad_log_prob_fn=target_grad_log_prob,
            init_position=position, num_steps=num_steps, schedule_type=schedule_type or "constant",
            target_accept=target_accept, max_iter_step=kwargs.get("max_iter_step", 1000),
            inv_mass_matrix=inv_mass_matrix, init_step_size=final_step_size, gamma_coarse_values=None,
            gamma_samples_per_eval=150, _chain_offset=chain_offset, verbose=verbose)
        gamma, steepness, final_step_size = tuned_gamma, tuned_steepness, tuned_step
        _log(verbose, f"  Refined friction parameters:\n    gamma={tuned_gamma:.5f}\n    steepness={tuned_steepness:.5f}"
                      f"\n    step_size={final_step_size:.5f} (re-tuned)")

    info = {
        "elapsed_time": time.time() - start_time,
        "final_step_size": final_step_size,
        "inv_mass_matrix": inv_mass_matrix,
        "mass_matrix_learned": learn_mass_matrix,
        "num_da_updates": n_batches_total,
        "grad_evals_local": grad_evals,
    }
    if grahmc:
        info["gamma"] = float(gamma) if gamma is not None else 1.0
        info["steepness"] = float(steepness) if steepness is not None else 5.0
        if tune_history is not None:
            info["friction_tuning"] = tune_history
    return final_step_size, inv_mass_matrix, position, info
