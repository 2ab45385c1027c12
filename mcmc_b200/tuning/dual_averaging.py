"""Dual-averaging step-size tuning, the reference's API (tuning/dual_averaging.py) over the
CUDA samplers.

State machines (da_*, joint_da_*) are host scalars, restated from
tuning/dual_averaging.py:806-887 and :608-666.  The tuner loops call the samplers once per
DA iteration with burn_in=0 exactly as the reference does (:76,194,320,770) but never store
the samples (MCMC_FLAG_NO_SAMPLES): only the fused in-kernel accept counter, all-reduced over
the chain shards (mcmc_b200/dist.py), comes back to the host.
"""
from __future__ import annotations

import math
from typing import Dict, NamedTuple, Tuple

import numpy as np
import torch

from .. import _abi, dist
from ..benchmarks.targets import resolve_log_prob
from ..random import split
from ..runtime import run_sampler, sum_f64
from ..samplers.GRAHMC import _schedule_id, get_friction_schedule


# ------------------------------------------------------------------ step-wise dual averaging
class DualAveragingState(NamedTuple):
    """tuning/dual_averaging.py:806-815."""
    log_step: float
    log_step_bar: float
    H_bar: float
    mu: float
    count: int
    omega: float = 0.05
    t0: float = 10.0
    kappa: float = 0.75


def da_init(initial_step_size: float) -> DualAveragingState:
    """tuning/dual_averaging.py:818-827 (mu = log eps0, not log(10 eps0))."""
    log_step = math.log(float(initial_step_size))
    return DualAveragingState(log_step=log_step, log_step_bar=log_step, H_bar=0.0, mu=log_step, count=0)


def da_update(state: DualAveragingState, accept_stat: float, target_accept: float) -> DualAveragingState:
    """tuning/dual_averaging.py:830-862."""
    m = state.count + 1
    eta_m = 1.0 / (m + state.t0)
    H_bar = (1 - eta_m) * state.H_bar + eta_m * (target_accept - accept_stat)
    log_step = state.mu - (math.sqrt(m) / state.omega) * H_bar
    m_kappa = m ** (-state.kappa)
    if m == 1:
        log_step_bar = log_step
    else:
        log_step_bar = m_kappa * log_step + (1 - m_kappa) * state.log_step_bar
    return DualAveragingState(log_step=float(log_step), log_step_bar=float(log_step_bar), H_bar=float(H_bar),
                              mu=state.mu, count=m, omega=state.omega, t0=state.t0, kappa=state.kappa)


def da_reset(state: DualAveragingState) -> DualAveragingState:
    """tuning/dual_averaging.py:864-887: restart around the smoothed estimate."""
    current = state.log_step_bar if state.count > 0 else state.log_step
    return DualAveragingState(log_step=current, log_step_bar=current, H_bar=0.0, mu=current, count=0,
                              omega=state.omega, t0=state.t0, kappa=state.kappa)


class DeviceDualAveraging:
    """The same state machine (da_init / da_update / da_reset) held in a float64[8] CUDA tensor and advanced by
    one-thread kernels (mcmc_da_init / mcmc_da_update / mcmc_da_reset): the sampler reads epsilon from `step_dev`
    (element 5) and the accept statistic never visits the host, so a whole warmup is enqueued without a sync."""

    def __init__(self, initial_step_size: float, device, omega: float = 0.05, t0: float = 10.0, kappa: float = 0.75):
        import ctypes as C
        self._C = C
        self.omega, self.t0, self.kappa = float(omega), float(t0), float(kappa)
        self.state = torch.empty((8,), dtype=torch.float64, device=device)
        self._L = _abi.lib()
        _abi.check(self._L.mcmc_set_device(self.state.device.index))
        _abi.check(self._L.mcmc_da_init(C.c_void_p(self.state.data_ptr()), float(initial_step_size), self._stream()))

    def _stream(self):
        return self._C.c_void_p(torch.cuda.current_stream(self.state.device).cuda_stream)

    @property
    def step_dev(self) -> torch.Tensor:
        return self.state[5:6]

    def update(self, accept_sum: torch.Tensor, n_terms: float, target_accept: float) -> None:
        """accept_sum: 1-element float64 CUDA tensor (already summed over the ranks)."""
        a = accept_sum.reshape(-1)
        assert a.is_cuda and a.dtype == torch.float64
        _abi.check(self._L.mcmc_da_update(self._C.c_void_p(self.state.data_ptr()), self._C.c_void_p(a.data_ptr()),
                                          float(n_terms), float(target_accept), self.omega, self.t0, self.kappa,
                                          self._stream()))
        a.record_stream(torch.cuda.current_stream(self.state.device))

    def reset(self) -> None:
        _abi.check(self._L.mcmc_da_reset(self._C.c_void_p(self.state.data_ptr()), self._stream()))

    def to_host(self) -> DualAveragingState:
        """Synchronises: the host view of the state."""
        v = self.state.cpu().numpy()
        return DualAveragingState(log_step=float(v[0]), log_step_bar=float(v[1]), H_bar=float(v[2]), mu=float(v[3]),
                                  count=int(v[4]), omega=self.omega, t0=self.t0, kappa=self.kappa)


# ------------------------------------------------------------------ joint (step, gamma) dual averaging
class JointDualAveragingState(NamedTuple):
    """tuning/dual_averaging.py:608-617."""
    log_params: np.ndarray
    log_params_bar: np.ndarray
    H_bar: float
    mu: np.ndarray
    count: int
    omega: float = 0.05
    t0: float = 10.0
    kappa: float = 0.75


def joint_da_init(initial_params) -> JointDualAveragingState:
    lp = np.log(np.asarray(initial_params, dtype=np.float64))
    return JointDualAveragingState(log_params=lp, log_params_bar=lp, H_bar=0.0, mu=lp, count=0)


def joint_da_update(state: JointDualAveragingState, accept_stat: float, target_accept: float) -> JointDualAveragingState:
    """tuning/dual_averaging.py:630-666 (gamma clipped to [0.01, 20])."""
    m = state.count + 1
    eta_m = 1.0 / (m + state.t0)
    H_bar = (1 - eta_m) * state.H_bar + eta_m * (target_accept - accept_stat)
    lp = state.mu - (math.sqrt(m) / state.omega) * H_bar
    lp = np.array([lp[0], np.clip(lp[1], math.log(0.01), math.log(20.0))])
    m_kappa = m ** (-state.kappa)
    lpb = m_kappa * lp + (1 - m_kappa) * state.log_params_bar
    return JointDualAveragingState(log_params=lp, log_params_bar=lpb, H_bar=float(H_bar), mu=state.mu, count=m,
                                   omega=state.omega, t0=state.t0, kappa=state.kappa)


# ------------------------------------------------------------------ one tuning batch on the device
def sample_batch(sampler: int, log_prob_fn, position, key, num_samples: int, *, step_size: float, num_steps: int = 0,
                 inv_mass_matrix=None, gamma: float = 0.0, steepness: float = 1.0, schedule_id: int = 0,
                 max_tree_depth: int = 10, welford=None, welford_count: float = 0.0, store_samples: bool = False,
                 track_proposals: bool = False, chain_offset: int = 0, step_size_dev=None) -> dict:
    """One `*_run(..., burn_in=0)` call of a tuning loop.  Returns the sampler's output dict plus
    `accept_sum` (0-d float64 tensor: this shard's sum of the DA statistic) and `accept_terms`."""
    tag = resolve_log_prob(log_prob_fn)
    nuts = sampler == _abi.SAMPLER_NUTS
    out = run_sampler(sampler, tag, position, key=key, step_size=float(step_size), num_steps=int(num_steps),
                      num_samples=int(num_samples), burn_in=0, inv_mass_matrix=inv_mass_matrix, gamma=float(gamma),
                      steepness=float(steepness), schedule_id=int(schedule_id), max_tree_depth=int(max_tree_depth),
                      welford=welford, welford_count=welford_count, track_proposals=track_proposals,
                      store_samples=bool(store_samples or nuts), chain_offset=chain_offset, step_size_dev=step_size_dev)
    n_chains = out["final_q"].shape[0]
    if nuts:
        # NUTS adapts on the mean Metropolis acceptance probability of the trajectories
        # (adaptation.py:187; dual_averaging.py:328)
        out["accept_sum"] = sum_f64(out["mean_accept_probs"])
    else:
        out["accept_sum"] = out["counters"][0].to(torch.float64)   # exact integer sum over chains, fused in-kernel
    out["accept_terms"] = n_chains * int(num_samples)
    return out


def _tune_loop(run_one, init_value: float, target_accept: float, tolerance: float, max_iter: int, min_iter: int,
               patience: int, label: str, fmt_extra=None):
    """The convergence loop shared by dual_averaging_tune_{rwmh,hmc,nuts}
    (tuning/dual_averaging.py:46-128, 164-251, 290-382): gamma=0.05, t0=10, kappa=0.75, mu = log(initial),
    log_bar initialised to the initial value, converged when the smoothed value moves < tolerance for
    `patience` consecutive iterations after `min_iter`."""
    gamma, t0, kappa = 0.05, 10.0, 0.75
    log_v = math.log(init_value)
    mu = log_v
    log_bar = log_v
    H_bar = 0.0
    value = math.exp(log_v)
    prev_bar = value
    hist_v, hist_a, extra = [], [], []
    converged_count, converged_iter = 0, max_iter
    for m in range(1, max_iter + 1):
        alpha, ex = run_one(value)
        eta_m = 1.0 / (m + t0)
        H_bar = (1 - eta_m) * H_bar + eta_m * (target_accept - alpha)
        log_v = mu - (math.sqrt(m) / gamma) * H_bar
        m_kappa = m ** (-kappa)
        log_bar = m_kappa * log_v + (1 - m_kappa) * log_bar
        value = math.exp(log_v)
        cur_bar = math.exp(log_bar)
        hist_v.append(cur_bar)
        hist_a.append(alpha)
        extra.append(ex)
        if m >= min_iter:
            rel = abs(cur_bar - prev_bar) / (abs(prev_bar) + 1e-10)
            converged_count = converged_count + 1 if rel < tolerance else 0
            if converged_count >= patience:
                if dist.rank() == 0:
                    print(f"  Converged after {m} iterations: {label}={cur_bar:.4f}, accept={alpha:.3f}")
                converged_iter = m
                return cur_bar, hist_v, hist_a, extra, converged_iter
        prev_bar = cur_bar
        if m % 200 == 0 and dist.rank() == 0:
            print(f"  Tuning iteration {m}: {label}={cur_bar:.4f}, accept={alpha:.3f}")
    final = math.exp(log_bar)
    if dist.rank() == 0:
        print(f"  Reached max iterations ({max_iter}): {label}={final:.4f}")
    return final, hist_v, hist_a, extra, converged_iter


def dual_averaging_tune_rwmh(key, log_prob_fn, init_position, target_accept: float = 0.234, tolerance: float = 0.01,
                             max_iter: int = 2000, min_iter: int = 100, patience: int = 10, *,
                             _chain_offset: int = 0) -> Tuple[float, Dict]:
    """Tune the RWMH proposal scale (tuning/dual_averaging.py:16-128); start 2.38/sqrt(d), 100 samples per iteration."""
    d = init_position.shape[-1]
    state = {"pos": init_position, "key": key}

    def run_one(scale):
        state["key"], sub = split(state["key"])
        out = sample_batch(_abi.SAMPLER_RWMH, log_prob_fn, state["pos"], sub, 100, step_size=scale,
                           chain_offset=_chain_offset)
        state["pos"] = out["final_q"]
        return dist.mean_accept(out["accept_sum"], out["accept_terms"]), None

    v, hv, ha, _, ci = _tune_loop(run_one, 2.38 / math.sqrt(d), target_accept, tolerance, max_iter, min_iter, patience,
                                  "scale")
    return v, {"scale_history": hv, "accept_history": ha, "converged_iter": ci, "target_accept": target_accept}


def dual_averaging_tune_hmc(key, log_prob_fn, init_position, num_steps: int, target_accept: float = 0.65,
                            tolerance: float = 0.01, max_iter: int = 2000, min_iter: int = 100, patience: int = 10, *,
                            _chain_offset: int = 0) -> Tuple[float, Dict]:
    """Tune the HMC step size for fixed L (tuning/dual_averaging.py:131-251); start 0.5/sqrt(d)."""
    d = init_position.shape[-1]
    state = {"pos": init_position, "key": key}

    def run_one(eps):
        state["key"], sub = split(state["key"])
        out = sample_batch(_abi.SAMPLER_HMC, log_prob_fn, state["pos"], sub, 100, step_size=eps, num_steps=num_steps,
                           chain_offset=_chain_offset)
        state["pos"] = out["final_q"]
        return dist.mean_accept(out["accept_sum"], out["accept_terms"]), None

    v, hv, ha, _, ci = _tune_loop(run_one, 0.5 / math.sqrt(d), target_accept, tolerance, max_iter, min_iter, patience,
                                  "step_size")
    return v, {"step_size_history": hv, "accept_history": ha, "converged_iter": ci, "target_accept": target_accept,
               "num_steps": num_steps}


def dual_averaging_tune_nuts(key, log_prob_fn, init_position, max_tree_depth: int = 10, target_accept: float = 0.65,
                             tolerance: float = 0.01, max_iter: int = 2000, min_iter: int = 100, patience: int = 10, *,
                             _chain_offset: int = 0) -> Tuple[float, Dict]:
    """Tune the NUTS step size (tuning/dual_averaging.py:254-382); the DA statistic is the mean
    Metropolis acceptance probability of the trajectories."""
    d = init_position.shape[-1]
    state = {"pos": init_position, "key": key}

    def run_one(eps):
        state["key"], sub = split(state["key"])
        out = sample_batch(_abi.SAMPLER_NUTS, log_prob_fn, state["pos"], sub, 100, step_size=eps,
                           max_tree_depth=max_tree_depth, chain_offset=_chain_offset)
        state["pos"] = out["final_q"]
        depth_sum = out["counters"][2].to(torch.float64)
        both = dist.global_mean(torch.stack([out["accept_sum"], depth_sum]), out["accept_terms"])
        return float(both[0].item()), float(both[1].item())

    v, hv, ha, depth, ci = _tune_loop(run_one, 0.5 / math.sqrt(d), target_accept, tolerance, max_iter, min_iter,
                                      patience, "step_size")
    return v, {"step_size_history": hv, "accept_history": ha, "tree_depth_history": depth, "converged_iter": ci,
               "target_accept": target_accept, "max_tree_depth": max_tree_depth}


def joint_tune_grahmc(key, log_prob_fn, grad_log_prob_fn, init_position, num_steps: int,
                      schedule_type: str = "constant", target_accept: float = 0.65, max_iter: int = 1000,
                      inv_mass_matrix=None, current_step_size: float = None, fixed_steepness: float = 10.0, *,
                      _chain_offset: int = 0) -> Tuple[float, float, float, Dict]:
    """Joint (step_size, gamma) dual averaging for GRAHMC (tuning/dual_averaging.py:668-804).
    grad_log_prob_fn is unused (kept for API compatibility, as in the reference)."""
    sched_id = _schedule_id(get_friction_schedule(schedule_type))
    n_dim = init_position.shape[-1]
    init_step = 0.5 / math.sqrt(n_dim) if current_step_size is None else float(current_step_size)
    state = joint_da_init(np.array([init_step, 1.0]))
    history = {"step_size": [], "gamma": [], "accept_rate": []}
    pos = init_position
    for m in range(1, max_iter + 1):
        key, sub = split(key)
        cur = np.exp(state.log_params)
        cur_step = float(cur[0])
        cur_gamma = max(0.001, min(float(cur[1]), 50.0))
        out = sample_batch(_abi.SAMPLER_GRAHMC, log_prob_fn, pos, sub, 50, step_size=cur_step, num_steps=num_steps,
                           gamma=cur_gamma, steepness=fixed_steepness, schedule_id=sched_id,
                           inv_mass_matrix=inv_mass_matrix, chain_offset=_chain_offset)
        pos = out["final_q"]
        alpha = dist.mean_accept(out["accept_sum"], out["accept_terms"])
        state = joint_da_update(state, alpha, target_accept)
        smooth = np.exp(state.log_params_bar)
        history["step_size"].append(float(smooth[0]))
        history["gamma"].append(float(smooth[1]))
        history["accept_rate"].append(alpha)
        if m % 100 == 0 and dist.rank() == 0:
            print(f"      Iter {m}: step={smooth[0]:.4f}, gamma={smooth[1]:.4f}, accept={alpha:.3f}")
    final = np.exp(state.log_params_bar)
    return float(final[0]), float(final[1]), fixed_steepness, history
