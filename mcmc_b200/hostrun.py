"""NumPy-level driver of the C ABI's host-pointer entry (mcmc_run_host).

This is the call a non-torch consumer makes (ctypes / cgo / a jax CPU custom
call): host arrays in, host arrays out, H2D and D2H copies owned by the
library.  The torch-tensor path used by the reference-signature wrappers is
mcmc_b200/runtime.py (device pointers, caller's stream).
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _abi


def make_params(target_id, target_params, n_chains, n_dim, num_samples, burn_in=0, num_steps=0, step_size=0.0,
                gamma=0.0, steepness=1.0, schedule_id=0, rng_mode=_abi.RNG_PHILOX, seed=(0, 0), chain_offset=0,
                flags=0, max_tree_depth=10, delta_max=1000.0) -> _abi.McmcParams:
    p = _abi.McmcParams()
    p.abi_version = _abi.ABI_VERSION
    p.target_id = int(target_id)
    p.schedule_id = int(schedule_id)
    p.rng_mode = int(rng_mode)
    tp = list(target_params) + [0.0] * (8 - len(target_params))
    for i in range(8):
        p.target_params[i] = float(tp[i])
    p.n_chains, p.num_samples, p.burn_in = int(n_chains), int(num_samples), int(burn_in)
    p.chain_offset = int(chain_offset)
    p.n_dim, p.num_steps, p.max_tree_depth, p.flags = int(n_dim), int(num_steps), int(max_tree_depth), int(flags)
    p.step_size, p.gamma, p.steepness, p.delta_max = float(step_size), float(gamma), float(steepness), float(delta_max)
    p.seed[0], p.seed[1] = int(seed[0]) & 0xFFFFFFFF, int(seed[1]) & 0xFFFFFFFF
    return p


def _ptr(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def run_host(sampler: int, p: _abi.McmcParams, init_q, inv_mass=None, cov_inv=None, z=None, u=None, dir_u=None,
             prop_u=None, welford=None, welford_count=0.0, device: int = 0, variant: str | None = None,
             outputs: dict | None = None) -> dict:
    """Run one sampler call with host buffers; returns a dict of NumPy arrays.

    `outputs` may pre-supply (e.g. pinned) arrays for any output field to avoid
    re-allocation between calls.
    """
    L = _abi.lib(variant)
    Cn, D, S = int(p.n_chains), int(p.n_dim), int(p.num_samples)
    f64 = lambda a: None if a is None else np.ascontiguousarray(a, dtype=np.float64)
    init_q = f64(init_q)
    if init_q.shape != (Cn, D):
        raise ValueError(f"init_q shape {init_q.shape} != ({Cn}, {D})")
    ins = dict(init_q=init_q, inv_mass=f64(inv_mass), cov_inv=f64(cov_inv), z_inject=f64(z), u_inject=f64(u),
               nuts_dir_u=f64(dir_u), nuts_prop_u=f64(prop_u))
    track = bool(p.flags & _abi.FLAG_TRACK_PROPOSALS) and sampler in (_abi.SAMPLER_HMC, _abi.SAMPLER_GRAHMC)
    store = not (p.flags & _abi.FLAG_NO_SAMPLES)
    nuts = sampler == _abi.SAMPLER_NUTS
    rw = sampler == _abi.SAMPLER_RWMH
    spec = dict(
        samples=((S, Cn, D), np.float64, store), log_probs=((S, Cn), np.float64, store),
        accept_count=((Cn,), np.int32, True), accept_rate=((Cn,), np.float32, True),
        final_q=((Cn, D), np.float64, True), final_lp=((Cn,), np.float64, True),
        final_g=((Cn, D), np.float64, not rw),
        pre_q=((S, Cn, D), np.float64, track), pre_lp=((S, Cn), np.float64, track),
        prop_q=((S, Cn, D), np.float64, track), prop_lp=((S, Cn), np.float64, track),
        delta_h=((S, Cn), np.float64, track),
        tree_depths=((S, Cn), np.int32, nuts), mean_accept_probs=((S, Cn), np.float64, nuts),
        counters=((4,), np.uint64, True),
    )
    out = {}
    for k, (shape, dt, want) in spec.items():
        if not want:
            continue
        if outputs is not None and k in outputs:
            a = outputs[k]
            assert a.shape == shape and a.dtype == dt and a.flags.c_contiguous
            out[k] = a
        else:
            out[k] = np.zeros(shape, dtype=dt)
    if welford is not None:
        out["welford_mean"] = np.array(welford[0], dtype=np.float64, copy=True)
        out["welford_m2"] = np.array(welford[1], dtype=np.float64, copy=True)
    b = _abi.McmcBuffers()
    for k, v in {**ins, **out}.items():
        setattr(b, k, _ptr(v))
    b.welford_count = float(welford_count)
    _abi.check(L.mcmc_run_host(int(sampler), C.byref(p), C.byref(b), int(device)), variant)
    return out
