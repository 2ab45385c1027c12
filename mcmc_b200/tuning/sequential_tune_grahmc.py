"""GRAHMC phase-3 tuning: step size by dual averaging, then friction amplitude by an ESJD
grid search.

The reference imports `tuning.sequential_tune_grahmc` (tuning/adaptation.py:282) but the module
is ABSENT from the reference tree.  This file is therefore not a restatement of
reference code: it implements what the reference *specifies* about it —
  * the call site, tuning/adaptation.py:287-300: arguments, `max_iter_step=1000` "warmup
    iterations per gamma", `gamma_coarse_values=None` -> "default [0.01, 0.1, 0.5, 1.0, 2.0, 5.0]",
    `gamma_samples_per_eval=150` "samples for ESJD measurement per gamma", `init_step_size` = the
    phase-2 step size as the starting point, returns (step, gamma, steepness, history);
  * the description in tuning/dual_averaging.py:683-708: "Phase 1: tune step_size with
    conservative gamma=0.5 (dual averaging); Phase 2: tune gamma via ESJD coarse-to-fine grid
    search".
ESJD = mean over chains and iterations of |q_{t+1} - q_t|^2 in the metric M (whitened by the
learned inv_mass_matrix), reduced on the device and all-reduced over the chain shards.
"""
from __future__ import annotations

import math
from typing import Dict, Optional, Sequence, Tuple

import torch

from .. import _abi, dist
from ..random import split
from ..samplers.GRAHMC import _schedule_id, get_friction_schedule
from .dual_averaging import da_init, da_update, sample_batch

DEFAULT_GAMMA_COARSE = (0.01, 0.1, 0.5, 1.0, 2.0, 5.0)


def _tune_step(key, log_prob_fn, position, num_steps, gamma, steepness, sched_id, inv_mass_matrix, init_step,
               target_accept, max_iter, chain_offset, batch=50):
    """Dual averaging on the step size at fixed friction; max_iter sampler iterations in batches."""
    state = da_init(init_step)
    for _ in range(max(1, max_iter // batch)):
        key, sub = split(key)
        out = sample_batch(_abi.SAMPLER_GRAHMC, log_prob_fn, position, sub, batch, step_size=math.exp(state.log_step),
                           num_steps=num_steps, gamma=gamma, steepness=steepness, schedule_id=sched_id,
                           inv_mass_matrix=inv_mass_matrix, chain_offset=chain_offset)
        position = out["final_q"]
        state = da_update(state, dist.mean_accept(out["accept_sum"], out["accept_terms"]), target_accept)
    return float(math.exp(state.log_step_bar)), position, key


def _esjd(key, log_prob_fn, position, num_steps, step, gamma, steepness, sched_id, inv_mass_matrix, n_samples,
          chain_offset):
    key, sub = split(key)
    out = sample_batch(_abi.SAMPLER_GRAHMC, log_prob_fn, position, sub, n_samples, step_size=step, num_steps=num_steps,
                       gamma=gamma, steepness=steepness, schedule_id=sched_id, inv_mass_matrix=inv_mass_matrix,
                       store_samples=True, chain_offset=chain_offset)
    s = out["samples"]
    prev = torch.cat([position.reshape(1, *s.shape[1:]).to(s.device), s[:-1]], dim=0)
    d2 = (s - prev) ** 2
    if inv_mass_matrix is not None:
        d2 = d2 / inv_mass_matrix.to(s.device)
    local = d2.sum()
    esjd = float(dist.global_mean(local.reshape(1), s.shape[0] * s.shape[1]).item())
    acc = dist.mean_accept(out["accept_sum"], out["accept_terms"])
    return esjd, acc, out["final_q"], key


def sequential_tune_grahmc(key, log_prob_fn, grad_log_prob_fn, init_position, num_steps: int,
                           schedule_type: str = "constant", target_accept: float = 0.65, max_iter_step: int = 1000,
                           inv_mass_matrix=None, init_step_size: Optional[float] = None,
                           gamma_coarse_values: Optional[Sequence[float]] = None, gamma_samples_per_eval: int = 150,
                           *, _chain_offset: int = 0, verbose: bool = True) -> Tuple[float, float, float, Dict]:
    sched_id = _schedule_id(get_friction_schedule(schedule_type))
    steepness = 0.5 if schedule_type == "tanh" else 2.0      # the warmup's own defaults (adaptation.py:128)
    n_dim = init_position.shape[-1]
    step0 = float(init_step_size) if init_step_size is not None else 0.5 / math.sqrt(n_dim)
    say = (lambda *a: print(*a)) if (verbose and dist.rank() == 0) else (lambda *a: None)
    from ..runtime import to_device, _device
    position = to_device(init_position, _device(init_position.device if getattr(init_position, "is_cuda", False) else None))
    if inv_mass_matrix is not None:
        inv_mass_matrix = to_device(inv_mass_matrix, position.device)

    # phase 1: step size at the conservative friction gamma = 0.5
    step, position, key = _tune_step(key, log_prob_fn, position, num_steps, 0.5, steepness, sched_id, inv_mass_matrix,
                                     step0, target_accept, max_iter_step, _chain_offset)
    say(f"    [seq-tune] step_size={step:.5f} at gamma=0.5")
    history = {"coarse": [], "fine": [], "step_size_phase1": step}

    # phase 2: ESJD over the coarse grid, then a geometric refinement between the best point's neighbours
    coarse = list(gamma_coarse_values) if gamma_coarse_values is not None else list(DEFAULT_GAMMA_COARSE)
    best = (-1.0, coarse[0])
    for g in coarse:
        e, a, position, key = _esjd(key, log_prob_fn, position, num_steps, step, g, steepness, sched_id,
                                    inv_mass_matrix, gamma_samples_per_eval, _chain_offset)
        history["coarse"].append({"gamma": g, "esjd": e, "accept": a})
        say(f"    [seq-tune] gamma={g:<6g} ESJD={e:.5f} accept={a:.3f}")
        if e > best[0]:
            best = (e, g)
    i = coarse.index(best[1])
    lo = coarse[max(i - 1, 0)]
    hi = coarse[min(i + 1, len(coarse) - 1)]
    fine = [math.exp(math.log(lo) + (math.log(hi) - math.log(lo)) * k / 6.0) for k in range(1, 6)] if hi > lo else []
    for g in fine:
        if any(abs(g - c) < 1e-12 for c in coarse):
            continue
        e, a, position, key = _esjd(key, log_prob_fn, position, num_steps, step, g, steepness, sched_id,
                                    inv_mass_matrix, gamma_samples_per_eval, _chain_offset)
        history["fine"].append({"gamma": g, "esjd": e, "accept": a})
        if e > best[0]:
            best = (e, g)
    gamma = float(best[1])

    # re-tune the step size at the chosen friction
    step, position, key = _tune_step(key, log_prob_fn, position, num_steps, gamma, steepness, sched_id,
                                     inv_mass_matrix, step, target_accept, max_iter_step, _chain_offset)
    history.update({"gamma": gamma, "esjd": best[0], "step_size": step, "steepness": steepness,
                    "final_position": position})
    say(f"    [seq-tune] chosen gamma={gamma:.4f} (ESJD {best[0]:.5f}), step_size={step:.5f}")
    return step, gamma, steepness, history
