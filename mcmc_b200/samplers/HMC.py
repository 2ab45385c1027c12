"""hmc_run / hmc_init with the reference's signatures (samplers/HMC.py:61-82, 223-314).

The whole burn-in + sampling loop runs as one CUDA kernel launch
(mcmc_b200/csrc/hmc_kernel.cuh); results are CUDA float64 tensors with the
reference's shapes.
"""
from __future__ import annotations

from typing import NamedTuple, Optional

import torch

from .. import _abi
from ..benchmarks.targets import resolve_log_prob
from ..random import split
from ..runtime import ensure_batched, run_sampler, target_eval, to_device, _device
from ..runtime import leapfrog as _leapfrog_device


class HMCState(NamedTuple):
    """samplers/HMC.py:23-35."""
    position: torch.Tensor       # (n_chains, n_dim) float64
    log_prob: torch.Tensor       # (n_chains,) float64
    grad_log_prob: torch.Tensor  # (n_chains, n_dim) float64
    accept_count: torch.Tensor   # (n_chains,) int32


def hmc_init(init_position, log_prob_fn) -> HMCState:
    tag = resolve_log_prob(log_prob_fn)
    pos, _ = ensure_batched(to_device(init_position, _device(getattr(init_position, "device", None)
                                                             if getattr(init_position, "is_cuda", False) else None)))
    lp, g = target_eval(tag, pos, want_grad=True)
    return HMCState(pos, lp, g, torch.zeros(pos.shape[0], dtype=torch.int32, device=pos.device))


def leapfrog(position, momentum, step_size: float, log_prob, grad_log_prob, log_prob_fn, num_steps: int, inv_mass_matrix,
             *, _variant: Optional[str] = None):
    """Leapfrog integration for every chain (samplers/HMC.py:85-134), one kernel launch (csrc/leapfrog_kernel.cuh).
    Returns (final_position, final_momentum, final_grad_log_prob, final_log_prob) — the reference's order."""
    tag = resolve_log_prob(log_prob_fn)
    q, p, lp, g = _leapfrog_device(tag, position, momentum, log_prob, grad_log_prob, step_size, num_steps, inv_mass_matrix,
                                   variant=_variant)
    return q, p, g, lp


def _pack(out, state_cls, track):
    state = state_cls(out["final_q"], out["final_lp"], out["final_g"], out["accept_count"])
    base = (out["samples"], out["log_probs"], out["accept_rate"], state)
    if track:
        return base + (out["pre_q"], out["pre_lp"], out["prop_q"], out["prop_lp"], out["delta_h"])
    return base


def hmc_run(key, log_prob_fn, init_position, step_size: float, num_steps: int, num_samples: int, burn_in: int = 0,
            inv_mass_matrix=None, track_proposals: bool = False, *, _inject: Optional[dict] = None,
            _variant: Optional[str] = None, _chain_offset: int = 0):
    """Run HMC with parallel chains.  Returns (samples (S,C,D), log_probs (S,C), accept_rate (C,) f32,
    final_state) and, with track_proposals, + (pre_positions, pre_lps, prop_positions, prop_lps, deltas_H)."""
    tag = resolve_log_prob(log_prob_fn)
    out = run_sampler(_abi.SAMPLER_HMC, tag, init_position, key=key, step_size=float(step_size),
                      num_steps=int(num_steps), num_samples=int(num_samples), burn_in=int(burn_in),
                      inv_mass_matrix=inv_mass_matrix, track_proposals=bool(track_proposals), inject=_inject,
                      variant=_variant, chain_offset=_chain_offset)
    return _pack(out, HMCState, track_proposals)


def hmc_step(state: HMCState, log_prob_fn, step_size: float, num_steps: int, key, inv_mass_matrix=None,
             return_proposal: bool = False, *, _inject: Optional[dict] = None, _variant: Optional[str] = None,
             _chain_offset: int = 0):
    """One HMC transition from `state` (samplers/HMC.py:138-219): (next_key, new_state) and, with return_proposal,
    + (proposal_pos, proposal_log_prob, delta_H).  As in the reference the key is split once (`key, step_key`) and the
    step draws from step_key; the launch is a one-iteration hmc_run, which re-derives the cached log-prob and gradient
    from state.position (the same values bit for bit for a state this package produced)."""
    tag = resolve_log_prob(log_prob_fn)
    key, step_key = split(key)
    out = run_sampler(_abi.SAMPLER_HMC, tag, state.position, key=step_key, step_size=float(step_size),
                      num_steps=int(num_steps), num_samples=1, burn_in=0, inv_mass_matrix=inv_mass_matrix,
                      track_proposals=bool(return_proposal), inject=_inject, variant=_variant, chain_offset=_chain_offset)
    new_state = type(state)(out["final_q"], out["final_lp"], out["final_g"], state.accept_count + out["accept_count"])
    if return_proposal:
        return key, new_state, out["prop_q"][0], out["prop_lp"][0], out["delta_h"][0]
    return key, new_state
