"""rahmc_run / rahmc_init and the friction-schedule registry with the reference's signatures
(samplers/GRAHMC.py:69-122, 127-148, 353-456).

The five schedule callables exist for API compatibility (identity against
FRICTION_SCHEDULES selects the kernel's schedule id; calling one evaluates
gamma(t) on the host exactly as the reference's formula).  The kernel builds the
per-step table s_i = exp(-gamma_i*eps/2) in its prologue.
"""
from __future__ import annotations

import math
from typing import Callable, Dict, NamedTuple, Optional

import torch

from .. import _abi
from ..benchmarks.targets import TaggedLogProb, resolve_log_prob
from ..random import split
from ..runtime import run_sampler
from ..runtime import leapfrog as _leapfrog_device
from .HMC import _pack, hmc_init


# N(0, I) with its normaliser, for any dimension (samplers/GRAHMC.py:42-46): a tagged log_prob_fn like the factories'
standard_normal_log_prob = TaggedLogProb("standard_normal", 0, [1.0])


class RAHMCState(NamedTuple):
    """samplers/GRAHMC.py:49-53."""
    position: torch.Tensor
    log_prob: torch.Tensor
    grad_log_prob: torch.Tensor
    accept_count: torch.Tensor


def constant_schedule(t, T, gamma, steepness=None):
    return -gamma if t < T / 2 else +gamma


def tanh_schedule(t, T, gamma_max, steepness=5.0):
    return gamma_max * math.tanh(steepness * (2.0 * t / T - 1.0))


def sigmoid_schedule(t, T, gamma_max, steepness=10.0):
    return gamma_max * (2.0 / (1.0 + math.exp(-(steepness * (t / T - 0.5)))) - 1.0)


def linear_schedule(t, T, gamma_max, steepness=None):
    return -gamma_max + (2.0 * gamma_max * t / T)


def sine_schedule(t, T, gamma_max, steepness=None):
    return gamma_max * math.sin(math.pi * (t / T - 0.5))


def constant_schedule_default(t, T):
    """Constant schedule with gamma = 1 (samplers/GRAHMC.py:108-110)."""
    return constant_schedule(t, T, gamma=1.0, steepness=None)


FRICTION_SCHEDULES: Dict[str, Callable] = {
    "constant": constant_schedule,
    "tanh": tanh_schedule,
    "sigmoid": sigmoid_schedule,
    "linear": linear_schedule,
    "sine": sine_schedule,
}


def get_friction_schedule(schedule_type: str) -> Callable:
    return FRICTION_SCHEDULES[schedule_type]  # KeyError for unknown names, like GRAHMC.py:122


def _schedule_id(friction_schedule) -> int:
    if friction_schedule is None:
        return _abi.SCHEDULE_IDS["constant"]  # GRAHMC.py:395-396
    if isinstance(friction_schedule, str):
        get_friction_schedule(friction_schedule)
        return _abi.SCHEDULE_IDS[friction_schedule]
    for name, fn in FRICTION_SCHEDULES.items():
        if fn is friction_schedule:
            return _abi.SCHEDULE_IDS[name]
    raise NotImplementedError(
        "friction_schedule must be one of mcmc_b200.samplers.GRAHMC.FRICTION_SCHEDULES (or its name); "
        "arbitrary Python schedules cannot be compiled into the kernel")


def _conformal_leapfrog_step(position, momentum, step_size: float, gamma: float, log_prob, grad_log_prob, log_prob_fn,
                             inv_mass_matrix, *, _variant: Optional[str] = None):
    """One conformal leapfrog step with friction `gamma` (samplers/GRAHMC.py:152-211): (q, p, log_prob, grad_log_prob).
    Runs as a one-step trajectory of the constant schedule, whose first half gives -gamma_max (GRAHMC.py:69-72)."""
    tag = resolve_log_prob(log_prob_fn)
    return _leapfrog_device(tag, position, momentum, log_prob, grad_log_prob, step_size, 1, inv_mass_matrix, friction=True,
                            gamma=-float(gamma), steepness=1.0, schedule_id=_abi.SCHEDULE_IDS["constant"], variant=_variant)


def _trajectory_with_schedule(position, momentum, step_size: float, gamma_max: float, steepness: float, log_prob,
                              grad_log_prob, num_steps: int, log_prob_fn, friction_schedule, inv_mass_matrix, *,
                              _variant: Optional[str] = None):
    """A full leapfrog trajectory with time-dependent friction (samplers/GRAHMC.py:215-261), one kernel launch.
    Returns (position, momentum, log_prob, grad_log_prob) — the reference's order for this function."""
    tag = resolve_log_prob(log_prob_fn)
    return _leapfrog_device(tag, position, momentum, log_prob, grad_log_prob, step_size, num_steps, inv_mass_matrix,
                            friction=True, gamma=float(gamma_max), steepness=1.0 if steepness is None else float(steepness),
                            schedule_id=_schedule_id(friction_schedule), variant=_variant)


def rahmc_init(init_position, log_prob_fn) -> RAHMCState:
    return RAHMCState(*hmc_init(init_position, log_prob_fn))


def rahmc_run(key, log_prob_fn, init_position, step_size: float, num_steps: int, gamma: float, steepness: float,
              num_samples: int, burn_in: int = 0, inv_mass_matrix=None, friction_schedule=None,
              track_proposals: bool = False, *, _inject: Optional[dict] = None, _variant: Optional[str] = None,
              _chain_offset: int = 0):
    """Run GRAHMC/RAHMC with parallel chains (same return layout as hmc_run)."""
    tag = resolve_log_prob(log_prob_fn)
    out = run_sampler(_abi.SAMPLER_GRAHMC, tag, init_position, key=key, step_size=float(step_size),
                      num_steps=int(num_steps), gamma=float(gamma),
                      steepness=float(steepness if steepness is not None else 1.0),  # GRAHMC.py:409
                      schedule_id=_schedule_id(friction_schedule), num_samples=int(num_samples),
                      burn_in=int(burn_in), inv_mass_matrix=inv_mass_matrix, track_proposals=bool(track_proposals),
                      inject=_inject, variant=_variant, chain_offset=_chain_offset)
    return _pack(out, RAHMCState, track_proposals)


def rahmc_step(state: RAHMCState, step_size: float, num_steps: int, gamma_max: float, steepness: float, key, log_prob_fn,
               inv_mass_matrix=None, friction_schedule=None, return_proposal: bool = False, *,
               _inject: Optional[dict] = None, _variant: Optional[str] = None, _chain_offset: int = 0):
    """One GRAHMC transition from `state` (samplers/GRAHMC.py:265-349; note the reference's argument order, which
    differs from hmc_step's): (next_key, new_state)[, proposal_pos, proposal_log_prob, delta_H]."""
    tag = resolve_log_prob(log_prob_fn)
    key, step_key = split(key)
    out = run_sampler(_abi.SAMPLER_GRAHMC, tag, state.position, key=step_key, step_size=float(step_size),
                      num_steps=int(num_steps), gamma=float(gamma_max),
                      steepness=float(steepness if steepness is not None else 1.0),
                      schedule_id=_schedule_id(friction_schedule), num_samples=1, burn_in=0,
                      inv_mass_matrix=inv_mass_matrix, track_proposals=bool(return_proposal), inject=_inject,
                      variant=_variant, chain_offset=_chain_offset)
    new_state = RAHMCState(out["final_q"], out["final_lp"], out["final_g"], state.accept_count + out["accept_count"])
    if return_proposal:
        return key, new_state, out["prop_q"][0], out["prop_lp"][0], out["delta_h"][0]
    return key, new_state
