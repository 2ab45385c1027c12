"""Device-pointer path of the C ABI driven from torch tensors.

torch is plumbing here: it owns device memory and the CUDA stream; every number
is produced by the kernels of libmcmc_b200.so.  The reference-signature
wrappers (mcmc_b200/samplers/*.py) and the tuning code call `run_sampler`.
"""
from __future__ import annotations

import ctypes as C
from typing import Optional

import numpy as np
import torch

from . import _abi
from .hostrun import make_params
from .random import as_key


def _device(device=None) -> torch.device:
    if device is not None:
        d = torch.device(device)
        if d.type != "cuda":
            raise RuntimeError("mcmc_b200 runs on CUDA devices only (there is no CPU fallback)")
        return d if d.index is not None else torch.device("cuda", torch.cuda.current_device())
    if not torch.cuda.is_available():
        raise RuntimeError("mcmc_b200 needs a CUDA device (there is no CPU fallback)")
    return torch.device("cuda", torch.cuda.current_device())


def to_device(x, device: torch.device, dtype=torch.float64) -> Optional[torch.Tensor]:
    """NumPy array / CPU tensor / CUDA tensor -> contiguous CUDA tensor of `dtype`."""
    if x is None:
        return None
    if isinstance(x, torch.Tensor):
        t = x
    else:
        t = torch.from_numpy(np.ascontiguousarray(x))
    if t.device != device:
        t = t.to(device, non_blocking=True)
    if t.dtype != dtype:
        t = t.to(dtype)
    return t.contiguous()


def _ptr(t: Optional[torch.Tensor]):
    return None if t is None else C.c_void_p(t.data_ptr())


def _stream_ptr(device):
    return C.c_void_p(torch.cuda.current_stream(device).cuda_stream)


def ensure_batched(init_position):
    """(n_dim,) -> (1, n_dim); (n_chains, n_dim) unchanged; else ValueError (HMC.py:38-57)."""
    nd = init_position.ndim if hasattr(init_position, "ndim") else np.asarray(init_position).ndim
    if nd == 1:
        return init_position[None, :], True
    if nd == 2:
        return init_position, False
    raise ValueError("Input must have shape (n_dim,) or (n_chains, n_dim).")


def run_sampler(sampler: int, tag, init_position, *, key=0, step_size=0.0, num_steps=0, num_samples=0, burn_in=0,
                inv_mass_matrix=None, gamma=0.0, steepness=1.0, schedule_id=0, track_proposals=False,
                max_tree_depth=10, delta_max=1000.0, inject: Optional[dict] = None, welford=None,
                welford_count=0.0, store_samples=True, chain_offset=0, device=None,
                variant: Optional[str] = None, step_size_dev: Optional[torch.Tensor] = None) -> dict:
    """One *_run call on the device.  Returns a dict of CUDA tensors (reference shapes/dtypes).

    inject: dict(z=(T,C,D), u=(T,C)[, dir_u=(T,C,max_tree_depth), prop_u=...]) switches the kernels from
    the Philox stream to caller-supplied randomness (the parity tests' hook).
    welford: (mean (C,D), m2 (C,D)) CUDA tensors updated IN PLACE with every stored sample.
    step_size_dev: 1-element float64 CUDA tensor the kernel reads epsilon from instead of `step_size` (device-resident
    dual averaging: tuning.dual_averaging.DeviceDualAveraging).
    """
    if isinstance(init_position, torch.Tensor) and init_position.is_cuda and device is None:
        device = init_position.device
    dev = _device(device)
    L = _abi.lib(variant)
    q0 = to_device(init_position, dev)
    q0, _ = ensure_batched(q0)
    q0 = q0.contiguous()
    Cn, D = q0.shape
    if tag.dim and D != tag.dim:
        raise ValueError(f"init_position has n_dim={D} but the target was built with dim={tag.dim}")
    S, B = int(num_samples), int(burn_in)
    seed = as_key(key)
    flags = 0
    hmc_like = sampler in (_abi.SAMPLER_HMC, _abi.SAMPLER_GRAHMC)
    if track_proposals and hmc_like:
        flags |= _abi.FLAG_TRACK_PROPOSALS
    if welford is not None:
        flags |= _abi.FLAG_ACCUM_WELFORD
    if not store_samples:
        flags |= _abi.FLAG_NO_SAMPLES
    p = make_params(tag.target_id, tag.target_params, Cn, D, S, B, num_steps, step_size, gamma, steepness,
                    schedule_id, _abi.RNG_INJECTED if inject else _abi.RNG_PHILOX, seed, chain_offset, flags,
                    max_tree_depth, delta_max)
    f64 = dict(dtype=torch.float64, device=dev)
    T = {}
    T["init_q"] = q0
    T["inv_mass"] = to_device(inv_mass_matrix, dev)
    if T["inv_mass"] is not None and tuple(T["inv_mass"].shape) != (D,):
        raise ValueError(f"inv_mass_matrix must have shape ({D},)")
    T["cov_inv"] = to_device(tag.cov_inv, dev)
    if inject:
        T["z_inject"] = to_device(inject["z"], dev)
        T["u_inject"] = to_device(inject["u"], dev)
        T["nuts_dir_u"] = to_device(inject.get("dir_u"), dev)
        T["nuts_prop_u"] = to_device(inject.get("prop_u"), dev)
    out = {}
    if store_samples:
        out["samples"] = torch.empty((S, Cn, D), **f64)
        out["log_probs"] = torch.empty((S, Cn), **f64)
    out["accept_count"] = torch.empty((Cn,), dtype=torch.int32, device=dev)
    out["accept_rate"] = torch.empty((Cn,), dtype=torch.float32, device=dev)
    out["final_q"] = torch.empty((Cn, D), **f64)
    out["final_lp"] = torch.empty((Cn,), **f64)
    if sampler != _abi.SAMPLER_RWMH:
        out["final_g"] = torch.empty((Cn, D), **f64)
    if flags & _abi.FLAG_TRACK_PROPOSALS:
        for k in ("pre_q", "prop_q"):
            out[k] = torch.empty((S, Cn, D), **f64)
        for k in ("pre_lp", "prop_lp", "delta_h"):
            out[k] = torch.empty((S, Cn), **f64)
    if sampler == _abi.SAMPLER_NUTS:
        out["tree_depths"] = torch.empty((S, Cn), dtype=torch.int32, device=dev)
        out["mean_accept_probs"] = torch.empty((S, Cn), **f64)
    out["counters"] = torch.empty((4,), dtype=torch.int64, device=dev)
    b = _abi.McmcBuffers()
    for k, v in {**T, **out}.items():
        setattr(b, k, _ptr(v))
    if welford is not None:
        wm, w2 = welford
        assert wm.is_cuda and w2.is_cuda and wm.dtype == torch.float64 and wm.is_contiguous() and w2.is_contiguous()
        assert tuple(wm.shape) == (Cn, D) and tuple(w2.shape) == (Cn, D)
        b.welford_mean, b.welford_m2 = _ptr(wm), _ptr(w2)
        b.welford_count = float(welford_count)
    if step_size_dev is not None:
        assert step_size_dev.is_cuda and step_size_dev.dtype == torch.float64 and step_size_dev.numel() >= 1
        b.step_size_dev = _ptr(step_size_dev)
        T["step_size_dev"] = step_size_dev
    ws_bytes = C.c_size_t(0)
    _abi.check(L.mcmc_workspace_bytes(int(sampler), C.byref(p), C.byref(ws_bytes)), variant)
    ws = None
    if ws_bytes.value:
        ws = torch.empty((ws_bytes.value,), dtype=torch.uint8, device=dev)
        b.workspace, b.workspace_bytes = _ptr(ws), ws_bytes.value
    fn = {_abi.SAMPLER_HMC: L.mcmc_hmc_run, _abi.SAMPLER_GRAHMC: L.mcmc_grahmc_run,
          _abi.SAMPLER_RWMH: L.mcmc_rwmh_run, _abi.SAMPLER_NUTS: L.mcmc_nuts_run}.get(sampler)
    if fn is None:
        raise ValueError(f"Unknown sampler: {sampler}")
    _abi.check(L.mcmc_set_device(dev.index), variant)
    _abi.check(fn(C.byref(p), C.byref(b), _stream_ptr(dev)), variant)
    # inputs / workspace are only read by the enqueued kernel: keep them alive until the stream passes it
    for t in list(T.values()) + [ws]:
        if t is not None:
            t.record_stream(torch.cuda.current_stream(dev))
    return out


def target_eval(tag, x, want_grad=True, device=None, variant=None):
    """vmap(value_and_grad(log_prob_fn))(x) on the device (HMC.py:73)."""
    if isinstance(x, torch.Tensor) and x.is_cuda and device is None:
        device = x.device
    dev = _device(device)
    L = _abi.lib(variant)
    xt = to_device(x, dev)
    single = xt.ndim == 1
    xt = xt.reshape(-1, xt.shape[-1]).contiguous()
    n, D = xt.shape
    if tag.dim and D != tag.dim:
        raise ValueError(f"x has n_dim={D} but the target was built with dim={tag.dim}")
    lp = torch.empty((n,), dtype=torch.float64, device=dev)
    g = torch.empty((n, D), dtype=torch.float64, device=dev) if want_grad else None
    p = make_params(tag.target_id, tag.target_params, n, D, 0)
    ci = to_device(tag.cov_inv, dev)
    _abi.check(L.mcmc_set_device(dev.index), variant)
    _abi.check(L.mcmc_target_eval(C.byref(p), _ptr(ci), _ptr(xt), n, _ptr(lp), _ptr(g), _stream_ptr(dev)), variant)
    if single:
        return lp[0], (g[0] if g is not None else None)
    return lp, g


def leapfrog(tag, position, momentum, log_prob, grad_log_prob, step_size, num_steps, inv_mass_matrix=None, *,
             friction=False, gamma=0.0, steepness=1.0, schedule_id=0, device=None, variant=None):
    """num_steps leapfrog steps for every chain on the device (mcmc_leapfrog: HMC.py:85-134; with friction GRAHMC.py:
    215-261).  Returns (position, momentum, log_prob, grad_log_prob) as new CUDA float64 tensors."""
    if isinstance(position, torch.Tensor) and position.is_cuda and device is None:
        device = position.device
    dev = _device(device)
    L = _abi.lib(variant)
    q = to_device(position, dev)
    q, _ = ensure_batched(q)
    q = q.contiguous()
    Cn, D = q.shape
    if tag.dim and D != tag.dim:
        raise ValueError(f"position has n_dim={D} but the target was built with dim={tag.dim}")
    mom = to_device(momentum, dev).reshape(Cn, D).contiguous()
    g = to_device(grad_log_prob, dev).reshape(Cn, D).contiguous()
    lp = to_device(log_prob, dev).reshape(Cn).contiguous()
    w = to_device(inv_mass_matrix, dev)
    if w is not None and tuple(w.shape) != (D,):
        raise ValueError(f"inv_mass_matrix must have shape ({D},)")
    p = make_params(tag.target_id, tag.target_params, Cn, D, 0, 0, int(num_steps), float(step_size), float(gamma),
                    float(steepness), int(schedule_id))
    qo, po, go, lo = torch.empty_like(q), torch.empty_like(mom), torch.empty_like(g), torch.empty_like(lp)
    ci = to_device(tag.cov_inv, dev)
    _abi.check(L.mcmc_set_device(dev.index), variant)
    _abi.check(L.mcmc_leapfrog(C.byref(p), int(bool(friction)), _ptr(ci), _ptr(w), _ptr(q), _ptr(mom), _ptr(lp), _ptr(g),
                               _ptr(qo), _ptr(po), _ptr(lo), _ptr(go), _stream_ptr(dev)), variant)
    return qo, po, lo, go


def philox_fill(key, n_chains: int, n_dim: int, it_begin: int, n_it: int, chain_offset: int = 0, device=None,
                variant=None):
    """The native random stream of iterations [it_begin, it_begin + n_it): (z (n_it, C, D), log_u (n_it, C)) CUDA
    tensors — exactly the normals and accept / slice log-uniforms the sampler kernels draw for these counters."""
    dev = _device(device)
    L = _abi.lib(variant)
    p = make_params(0, [], n_chains, n_dim, 0, seed=as_key(key), chain_offset=chain_offset)
    z = torch.empty((n_it, n_chains, n_dim), dtype=torch.float64, device=dev)
    lu = torch.empty((n_it, n_chains), dtype=torch.float64, device=dev)
    _abi.check(L.mcmc_set_device(dev.index), variant)
    _abi.check(L.mcmc_philox_fill(C.byref(p), int(it_begin), int(n_it), _ptr(z), _ptr(lu), _stream_ptr(dev)), variant)
    return z, lu


def friction_table(schedule_id: int, num_steps: int, step_size: float, gamma: float, steepness: float,
                   device=None, variant=None):
    dev = _device(device)
    L = _abi.lib(variant)
    s = torch.empty((num_steps,), dtype=torch.float64, device=dev)
    g = torch.empty((num_steps,), dtype=torch.float64, device=dev)
    _abi.check(L.mcmc_set_device(dev.index), variant)
    _abi.check(L.mcmc_friction_table(int(schedule_id), int(num_steps), float(step_size), float(gamma),
                                     float(steepness), _ptr(s), _ptr(g), _stream_ptr(dev)), variant)
    return s, g


_REDUCE_WS = {}


def _reduce_ws(dev, nbytes):
    ws = _REDUCE_WS.get(dev)
    if ws is None or ws.numel() < nbytes:
        ws = torch.empty((max(nbytes, 1 << 20),), dtype=torch.uint8, device=dev)
        _REDUCE_WS[dev] = ws
    return ws


def welford_sum_var(welford_m2: torch.Tensor, count: float, variant=None) -> torch.Tensor:
    """sum over this shard's chains of m2/(max(count,2)-1): (D,) CUDA tensor (adaptation.py:240-248)."""
    dev = welford_m2.device
    L = _abi.lib(variant)
    Cn, D = welford_m2.shape
    out = torch.empty((D,), dtype=torch.float64, device=dev)
    ws = _reduce_ws(dev, 256 * D * 8)
    _abi.check(L.mcmc_set_device(dev.index), variant)
    _abi.check(L.mcmc_welford_finalize(_ptr(welford_m2), float(count), Cn, D, _ptr(out), _ptr(ws), ws.numel(),
                                       _stream_ptr(dev)), variant)
    return out


def chain_moments(samples: torch.Tensor, welford_mean: torch.Tensor, welford_m2: torch.Tensor,
                  count_before: float = 0.0, variant=None) -> None:
    """Fold samples (S,C,D) into per-chain Welford states in place, as S sequential welford_update calls."""
    dev = samples.device
    L = _abi.lib(variant)
    S, Cn, D = samples.shape
    _abi.check(L.mcmc_set_device(dev.index), variant)
    _abi.check(L.mcmc_chain_moments(_ptr(samples.contiguous()), S, Cn, D, _ptr(welford_mean), _ptr(welford_m2),
                                    float(count_before), _stream_ptr(dev)), variant)


def sum_f64(x: torch.Tensor, variant=None) -> torch.Tensor:
    """Deterministic (fixed-order) sum of a float64 CUDA tensor -> 0-d CUDA tensor."""
    dev = x.device
    L = _abi.lib(variant)
    xf = x.contiguous().reshape(-1)
    out = torch.empty((1,), dtype=torch.float64, device=dev)
    ws = _reduce_ws(dev, 256 * 8)
    _abi.check(L.mcmc_set_device(dev.index), variant)
    _abi.check(L.mcmc_sum_f64(_ptr(xf), xf.numel(), _ptr(out), _ptr(ws), ws.numel(), _stream_ptr(dev)), variant)
    return out[0]
