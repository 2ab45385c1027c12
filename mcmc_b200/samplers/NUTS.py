"""nuts_run / nuts_init with the reference's signatures (samplers/NUTS.py:73-86, 446-514)."""
from __future__ import annotations

from typing import NamedTuple, Optional

import torch

from .. import _abi
from ..benchmarks.targets import resolve_log_prob
from ..random import split
from ..runtime import run_sampler
from .HMC import hmc_init


class NUTSState(NamedTuple):
    """samplers/NUTS.py:30-42."""
    position: torch.Tensor
    log_prob: torch.Tensor
    grad_log_prob: torch.Tensor
    accept_count: torch.Tensor


def nuts_init(init_position, log_prob_fn) -> NUTSState:
    return NUTSState(*hmc_init(init_position, log_prob_fn))


def nuts_run(key, log_prob_fn, init_position, step_size: float, num_samples: int, burn_in: int = 0,
             inv_mass_matrix=None, max_tree_depth: int = 10, delta_max: float = 1000.0, *,
             _inject: Optional[dict] = None, _variant: Optional[str] = None, _chain_offset: int = 0):
    """Run NUTS (the reference's iterative-doubling variant).  Returns (samples, log_probs, accept_rate,
    final_state, tree_depths (S,C) int32, mean_accept_probs (S,C))."""
    tag = resolve_log_prob(log_prob_fn)
    out = run_sampler(_abi.SAMPLER_NUTS, tag, init_position, key=key, step_size=float(step_size),
                      num_samples=int(num_samples), burn_in=int(burn_in), inv_mass_matrix=inv_mass_matrix,
                      max_tree_depth=int(max_tree_depth), delta_max=float(delta_max), inject=_inject,
                      variant=_variant, chain_offset=_chain_offset)
    state = NUTSState(out["final_q"], out["final_lp"], out["final_g"], out["accept_count"])
    return (out["samples"], out["log_probs"], out["accept_rate"], state, out["tree_depths"],
            out["mean_accept_probs"])


def nuts_step(state: NUTSState, log_prob_fn, step_size: float, key, inv_mass_matrix=None, max_tree_depth: int = 10,
              delta_max: float = 1000.0, *, _inject: Optional[dict] = None, _variant: Optional[str] = None,
              _chain_offset: int = 0):
    """One NUTS transition for all chains from `state` (samplers/NUTS.py:399-442):
    (next_key, new_state, tree_depths (C,), mean_accept_probs (C,))."""
    tag = resolve_log_prob(log_prob_fn)
    key, step_key = split(key)
    out = run_sampler(_abi.SAMPLER_NUTS, tag, state.position, key=step_key, step_size=float(step_size), num_samples=1,
                      burn_in=0, inv_mass_matrix=inv_mass_matrix, max_tree_depth=int(max_tree_depth),
                      delta_max=float(delta_max), inject=_inject, variant=_variant, chain_offset=_chain_offset)
    new_state = NUTSState(out["final_q"], out["final_lp"], out["final_g"], state.accept_count + out["accept_count"])
    return key, new_state, out["tree_depths"][0], out["mean_accept_probs"][0]
