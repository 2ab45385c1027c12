"""jax.ffi registration of the sampling kernels (north_star: "calls the kernels through a thin C-ABI registered as a
jax.ffi custom call").

Where `import jax` works (not in this image: no jaxlib wheel, no network), `register()` builds
mcmc_b200/csrc/mcmc_xla_ffi.cc against jaxlib's own XLA FFI header, registers its six handlers for platform "CUDA" and
the functions below become jit-compatible drop-ins with the reference's signatures, taking and returning jax arrays:

    hmc_run   samplers/HMC.py:223-314      rahmc_run  samplers/GRAHMC.py:353-456
    rwMH_run  samplers/RWMH.py:123-178     nuts_run   samplers/NUTS.py:446-514

Without jax every entry raises ImportError — there is no CPU fallback and no silent substitute; the torch-tensor
wrappers of mcmc_b200/samplers/ are the host API that runs here.  Static arguments (log_prob_fn's target id and
parameters, the friction schedule, L, S, burn-in, step size, the key) travel as custom-call attributes, arrays as
buffers; the sampler workspace comes from XLA's scratch allocator.
"""
from __future__ import annotations

import ctypes
from typing import NamedTuple

import numpy as np

from . import _abi
from .benchmarks.targets import resolve_log_prob
from .random import as_key

_TARGETS = {"mcmc_hmc_run": "McmcHmcRun", "mcmc_grahmc_run": "McmcGrahmcRun", "mcmc_rwmh_run": "McmcRwmhRun",
            "mcmc_nuts_run": "McmcNutsRun", "mcmc_hmc_run_tracked": "McmcHmcRunTracked",
            "mcmc_grahmc_run_tracked": "McmcGrahmcRunTracked"}
_STATE = {"registered": False}


def available() -> bool:
    try:
        import jax  # noqa: F401
        return hasattr(jax, "ffi")
    except Exception:
        return False


def _jax():
    try:
        import jax
        import jax.numpy as jnp
        return jax, jnp
    except Exception as e:  # pragma: no cover - jax is absent from this image
        raise ImportError("mcmc_b200.ffi needs jax + jaxlib (not installed here); use mcmc_b200.samplers (torch CUDA "
                          "tensors) instead — there is no CPU fallback") from e


def register() -> None:
    """Build the shim if needed and register the custom-call targets (idempotent)."""
    if _STATE["registered"]:
        return
    jax, _ = _jax()
    from .csrc import build as cbuild
    _abi.lib("fast")                               # libmcmc_b200.so must be loaded first (the shim links against it)
    path = cbuild.build_xla_ffi(verbose=True)
    if path is None:
        raise ImportError("the XLA FFI shim could not be built (jax.ffi.include_dir() unavailable)")
    shim = ctypes.CDLL(path)
    if shim.mcmc_xla_ffi_available() != 1:
        raise ImportError("mcmc_xla_ffi.cc was compiled without xla/ffi/api/ffi.h")
    for name, sym in _TARGETS.items():
        jax.ffi.register_ffi_target(name, jax.ffi.pycapsule(getattr(shim, sym)), platform="CUDA")
    _STATE["registered"] = True


class SamplerState(NamedTuple):
    """HMCState / RAHMCState / NUTSState (samplers/HMC.py:23-35); RWMState has no gradient (RWMH.py:20-30)."""
    position: object
    log_prob: object
    grad_log_prob: object
    accept_count: object


def _call(target, tag, init_position, key, *, step_size=0.0, num_steps=0, num_samples=0, burn_in=0, inv_mass_matrix=None,
          gamma=0.0, steepness=1.0, schedule_id=0, max_tree_depth=10, delta_max=1000.0, chain_offset=0, extra=()):
    jax, jnp = _jax()
    register()
    pos = jnp.asarray(init_position, dtype=jnp.float64)
    if pos.ndim == 1:                              # HMC.py:38-57
        pos = pos[None, :]
    if pos.ndim != 2:
        raise ValueError("Input must have shape (n_dim,) or (n_chains, n_dim).")
    C, D = pos.shape
    S = int(num_samples)
    f64, f32, i32 = jnp.float64, jnp.float32, jnp.int32
    sds = jax.ShapeDtypeStruct
    outs = [sds((S, C, D), f64), sds((S, C), f64), sds((C,), f32), sds((C, D), f64), sds((C,), f64), sds((C, D), f64),
            sds((C,), i32)]
    for kind in extra:
        outs.append({"scd": sds((S, C, D), f64), "sc": sds((S, C), f64), "sc_i32": sds((S, C), i32)}[kind])
    k = as_key(key)
    w = jnp.zeros((0,), f64) if inv_mass_matrix is None else jnp.asarray(inv_mass_matrix, f64)
    P = jnp.zeros((0, 0), f64) if tag.cov_inv is None else jnp.asarray(tag.cov_inv, f64)
    tp = np.zeros(8, dtype=np.float64)
    tp[:len(tag.target_params)] = tag.target_params
    return jax.ffi.ffi_call(target, tuple(outs))(
        pos, w, P, target_id=np.int32(tag.target_id), schedule_id=np.int32(schedule_id), num_steps=np.int32(num_steps),
        max_tree_depth=np.int32(max_tree_depth), num_samples=np.int64(S), burn_in=np.int64(burn_in),
        chain_offset=np.int64(chain_offset), step_size=np.float64(step_size), gamma=np.float64(gamma),
        steepness=np.float64(steepness), delta_max=np.float64(delta_max), seed0=np.int64(int(k[0])),
        seed1=np.int64(int(k[1])), target_params=tp)


def _pack(o, track):
    base = (o[0], o[1], o[2], SamplerState(o[3], o[4], o[5], o[6]))
    return base + tuple(o[7:12]) if track else base


def hmc_run(key, log_prob_fn, init_position, step_size, num_steps, num_samples, burn_in=0, inv_mass_matrix=None,
            track_proposals=False):
    tag = resolve_log_prob(log_prob_fn)
    o = _call("mcmc_hmc_run_tracked" if track_proposals else "mcmc_hmc_run", tag, init_position, key, step_size=step_size,
              num_steps=num_steps, num_samples=num_samples, burn_in=burn_in, inv_mass_matrix=inv_mass_matrix,
              extra=("scd", "sc", "scd", "sc", "sc") if track_proposals else ())
    return _pack(o, track_proposals)


def rahmc_run(key, log_prob_fn, init_position, step_size, num_steps, gamma, steepness, num_samples, burn_in=0,
              inv_mass_matrix=None, friction_schedule=None, track_proposals=False):
    from .samplers.GRAHMC import _schedule_id, constant_schedule
    tag = resolve_log_prob(log_prob_fn)
    sid = _schedule_id(friction_schedule if friction_schedule is not None else constant_schedule)   # GRAHMC.py:395-396
    o = _call("mcmc_grahmc_run_tracked" if track_proposals else "mcmc_grahmc_run", tag, init_position, key,
              step_size=step_size, num_steps=num_steps, num_samples=num_samples, burn_in=burn_in,
              inv_mass_matrix=inv_mass_matrix, gamma=gamma, steepness=1.0 if steepness is None else steepness,
              schedule_id=sid, extra=("scd", "sc", "scd", "sc", "sc") if track_proposals else ())
    return _pack(o, track_proposals)


def rwMH_run(key, log_prob_fn, init_position, num_samples, scale, burn_in=0):
    tag = resolve_log_prob(log_prob_fn)
    o = _call("mcmc_rwmh_run", tag, init_position, key, step_size=scale, num_samples=num_samples, burn_in=burn_in)
    return o[0], o[1], o[2], SamplerState(o[3], o[4], None, o[6])


def nuts_run(key, log_prob_fn, init_position, step_size, num_samples, burn_in=0, inv_mass_matrix=None,
             max_tree_depth=10, delta_max=1000.0):
    tag = resolve_log_prob(log_prob_fn)
    o = _call("mcmc_nuts_run", tag, init_position, key, step_size=step_size, num_samples=num_samples, burn_in=burn_in,
              inv_mass_matrix=inv_mass_matrix, max_tree_depth=max_tree_depth, delta_max=delta_max, extra=("sc_i32", "sc"))
    return o[0], o[1], o[2], SamplerState(o[3], o[4], o[5], o[6]), o[7], o[8]
