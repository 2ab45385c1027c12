"""rwMH_run / rwMH_init with the reference's signatures (samplers/RWMH.py:55-69, 123-178)."""
from __future__ import annotations

from typing import NamedTuple, Optional

import torch

from .. import _abi
from ..benchmarks.targets import resolve_log_prob
from ..random import split
from ..runtime import ensure_batched, run_sampler, target_eval, to_device, _device


class RWMState(NamedTuple):
    """samplers/RWMH.py:20-30."""
    position: torch.Tensor
    log_prob: torch.Tensor
    accept_count: torch.Tensor


def rwMH_init(init_position, log_prob_fn) -> RWMState:
    tag = resolve_log_prob(log_prob_fn)
    pos, _ = ensure_batched(to_device(init_position, _device(getattr(init_position, "device", None)
                                                             if getattr(init_position, "is_cuda", False) else None)))
    lp, _ = target_eval(tag, pos, want_grad=False)
    return RWMState(pos, lp, torch.zeros(pos.shape[0], dtype=torch.int32, device=pos.device))


def rwMH_run(key, log_prob_fn, init_position, num_samples: int, scale: float, burn_in: int = 0, *,
             _inject: Optional[dict] = None, _variant: Optional[str] = None, _chain_offset: int = 0):
    """Run random-walk Metropolis.  Returns (samples, log_probs, accept_rate, final_state)."""
    tag = resolve_log_prob(log_prob_fn)
    out = run_sampler(_abi.SAMPLER_RWMH, tag, init_position, key=key, step_size=float(scale),
                      num_samples=int(num_samples), burn_in=int(burn_in), inject=_inject, variant=_variant,
                      chain_offset=_chain_offset)
    state = RWMState(out["final_q"], out["final_lp"], out["accept_count"])
    return out["samples"], out["log_probs"], out["accept_rate"], state


def rwMH_step(state: RWMState, log_prob_fn, scale: float, key, *, _inject: Optional[dict] = None,
              _variant: Optional[str] = None, _chain_offset: int = 0):
    """One random-walk Metropolis transition from `state` (samplers/RWMH.py:73-119): (next_key, new_state)."""
    tag = resolve_log_prob(log_prob_fn)
    key, step_key = split(key)
    out = run_sampler(_abi.SAMPLER_RWMH, tag, state.position, key=step_key, step_size=float(scale), num_samples=1,
                      burn_in=0, inject=_inject, variant=_variant, chain_offset=_chain_offset)
    return key, RWMState(out["final_q"], out["final_lp"], state.accept_count + out["accept_count"])
