"""ctypes driver of oracle/liboracle.so (the C oracle).

TEST INFRASTRUCTURE — imported only by tests/, __graft_entry__.smoke() and
bench.py's cpu_baseline / --impl reference legs.  Never by mcmc_b200/.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

from mcmc_b200 import _abi

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "liboracle.so")
_LIB = None


def build(force: bool = False) -> str:
    src = [os.path.join(_HERE, f) for f in ("mcmc_oracle.c", "rng_common.h")]
    src.append(os.path.join(_HERE, "..", "include", "mcmc_b200.h"))
    stale = (not os.path.exists(_SO)) or any(
        os.path.exists(s) and os.path.getmtime(s) > os.path.getmtime(_SO) for s in src)
    if force or stale:
        subprocess.check_call(["make", "-C", _HERE, "-B", "liboracle.so"], stdout=subprocess.DEVNULL)
    return _SO


def build_native(out_path: str) -> str:
    """The same source compiled for speed on THIS host (-O3 -march=native, FMA contraction allowed): what bench.py's
    cpu_baseline / --impl reference legs time, so that the CPU number is not handicapped by the checker's strict flags.
    Must be built on the machine that runs it (-march=native), hence not part of build()."""
    cc = "/usr/bin/gcc" if os.path.exists("/usr/bin/gcc") else "gcc"
    subprocess.check_call([cc, "-O3", "-march=native", "-fPIC", "-std=c11", "-ffp-contract=fast", "-fopenmp", "-D_GNU_SOURCE",
                           "-shared", "-o", out_path, os.path.join(_HERE, "mcmc_oracle.c"), "-lm"])
    return out_path


def use_library(path: str) -> None:
    """Point this driver at another build of the oracle (bench.py: the -march=native one)."""
    global _LIB, _SO
    _SO = path
    _LIB = None


def lib() -> C.CDLL:
    global _LIB
    if _LIB is None:
        if not os.path.exists(_SO):
            build()
        L = C.CDLL(_SO)
        for n in ("oracle_hmc_run", "oracle_grahmc_run", "oracle_rwmh_run", "oracle_nuts_run"):
            getattr(L, n).argtypes = [C.POINTER(_abi.McmcParams), C.POINTER(_abi.McmcBuffers)]
        L.oracle_target_eval.argtypes = [C.POINTER(_abi.McmcParams), C.c_void_p, C.c_void_p, C.c_int64,
                                         C.c_void_p, C.c_void_p]
        L.oracle_leapfrog.argtypes = [C.POINTER(_abi.McmcParams), C.c_int] + [C.c_void_p] * 10
        L.oracle_friction_table.argtypes = [C.c_int, C.c_int, C.c_double, C.c_double, C.c_double,
                                            C.c_void_p, C.c_void_p]
        L.oracle_friction_table.restype = None
        L.oracle_welford_finalize.argtypes = [C.c_void_p, C.c_double, C.c_int64, C.c_int32, C.c_void_p]
        L.oracle_philox4x32_10.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
        L.oracle_philox4x32_10.restype = None
        _LIB = L
    return _LIB


def num_threads() -> int:
    return int(lib().oracle_num_threads())


def use_all_cores() -> int:
    """Size the OpenMP team to the cores this process may run on, whatever OMP_NUM_THREADS says (torchrun sets it to 1
    in every worker's environment); returns the team size."""
    try:
        n = len(os.sched_getaffinity(0))
    except AttributeError:
        n = os.cpu_count() or 1
    lib().oracle_set_num_threads(int(n))
    return num_threads()


def _ptr(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def make_params(target_id, target_params, C_, D, S, burn_in=0, L=0, step_size=0.0, gamma=0.0, steepness=1.0,
                schedule_id=0, rng_mode=_abi.RNG_INJECTED, seed=(0, 0), chain_offset=0, flags=0,
                max_tree_depth=10, delta_max=1000.0) -> _abi.McmcParams:
    p = _abi.McmcParams()
    p.abi_version = _abi.ABI_VERSION
    p.target_id = int(target_id)
    p.schedule_id = int(schedule_id)
    p.rng_mode = int(rng_mode)
    tp = list(target_params) + [0.0] * (8 - len(target_params))
    for i in range(8):
        p.target_params[i] = float(tp[i])
    p.n_chains, p.num_samples, p.burn_in, p.chain_offset = int(C_), int(S), int(burn_in), int(chain_offset)
    p.n_dim, p.num_steps, p.max_tree_depth, p.flags = int(D), int(L), int(max_tree_depth), int(flags)
    p.step_size, p.gamma, p.steepness, p.delta_max = float(step_size), float(gamma), float(steepness), float(delta_max)
    p.seed[0], p.seed[1] = int(seed[0]) & 0xFFFFFFFF, int(seed[1]) & 0xFFFFFFFF
    return p


_RUN = {_abi.SAMPLER_HMC: "oracle_hmc_run", _abi.SAMPLER_GRAHMC: "oracle_grahmc_run",
        _abi.SAMPLER_RWMH: "oracle_rwmh_run", _abi.SAMPLER_NUTS: "oracle_nuts_run"}


def run(sampler: int, p: _abi.McmcParams, init_q, inv_mass=None, cov_inv=None, z=None, u=None,
        dir_u=None, prop_u=None, welford=None, welford_count=0.0, store_samples=True) -> dict:
    """Run the C oracle; every output is a fresh NumPy array."""
    Cn, D, S = int(p.n_chains), int(p.n_dim), int(p.num_samples)
    f64 = lambda a: None if a is None else np.ascontiguousarray(a, dtype=np.float64)
    init_q = f64(init_q)
    assert init_q.shape == (Cn, D)
    ins = dict(init_q=init_q, inv_mass=f64(inv_mass), cov_inv=f64(cov_inv), z_inject=f64(z), u_inject=f64(u),
               nuts_dir_u=f64(dir_u), nuts_prop_u=f64(prop_u))
    track = bool(p.flags & _abi.FLAG_TRACK_PROPOSALS)
    nuts = sampler == _abi.SAMPLER_NUTS
    rw = sampler == _abi.SAMPLER_RWMH
    out = dict(
        samples=np.zeros((S, Cn, D)) if store_samples else None,
        log_probs=np.zeros((S, Cn)) if store_samples else None,
        accept_count=np.zeros(Cn, dtype=np.int32), accept_rate=np.zeros(Cn, dtype=np.float32),
        final_q=np.zeros((Cn, D)), final_lp=np.zeros(Cn), final_g=None if rw else np.zeros((Cn, D)),
        pre_q=np.zeros((S, Cn, D)) if track else None, pre_lp=np.zeros((S, Cn)) if track else None,
        prop_q=np.zeros((S, Cn, D)) if track else None, prop_lp=np.zeros((S, Cn)) if track else None,
        delta_h=np.zeros((S, Cn)) if track else None,
        tree_depths=np.zeros((S, Cn), dtype=np.int32) if nuts else None,
        mean_accept_probs=np.zeros((S, Cn)) if nuts else None,
        counters=np.zeros(4, dtype=np.uint64),
    )
    if welford is not None:
        out["welford_mean"] = np.array(welford[0], dtype=np.float64, copy=True)
        out["welford_m2"] = np.array(welford[1], dtype=np.float64, copy=True)
    b = _abi.McmcBuffers()
    for k, v in {**ins, **out}.items():
        setattr(b, k, _ptr(v))
    b.welford_count = float(welford_count)
    rc = getattr(lib(), _RUN[sampler])(C.byref(p), C.byref(b))
    if rc != 0:
        raise ValueError(f"oracle returned {rc}")
    return {k: v for k, v in out.items() if v is not None}


def target_eval(p: _abi.McmcParams, q, cov_inv=None, want_grad=True):
    q = np.ascontiguousarray(q, dtype=np.float64)
    n = q.shape[0]
    lp = np.zeros(n)
    g = np.zeros_like(q) if want_grad else None
    ci = None if cov_inv is None else np.ascontiguousarray(cov_inv, dtype=np.float64)
    lib().oracle_target_eval(C.byref(p), _ptr(ci), _ptr(q), n, _ptr(lp), _ptr(g))
    return lp, g


def leapfrog(p: _abi.McmcParams, q, mom, lp, grad, inv_mass=None, cov_inv=None, friction=False):
    """(q, momentum, log_prob, grad) after p.num_steps steps: HMC.py:85-134 / GRAHMC.py:215-261 (friction=True)."""
    q = np.ascontiguousarray(q, dtype=np.float64)
    mom = np.ascontiguousarray(mom, dtype=np.float64)
    grad = np.ascontiguousarray(grad, dtype=np.float64)
    lp = None if lp is None else np.ascontiguousarray(lp, dtype=np.float64)
    w = None if inv_mass is None else np.ascontiguousarray(inv_mass, dtype=np.float64)
    ci = None if cov_inv is None else np.ascontiguousarray(cov_inv, dtype=np.float64)
    qo, po, go, lo = np.empty_like(q), np.empty_like(mom), np.empty_like(grad), np.empty(q.shape[0])
    rc = lib().oracle_leapfrog(C.byref(p), int(bool(friction)), _ptr(ci), _ptr(w), _ptr(q), _ptr(mom), _ptr(lp), _ptr(grad),
                               _ptr(qo), _ptr(po), _ptr(lo), _ptr(go))
    if rc != 0:
        raise ValueError(f"oracle returned {rc}")
    return qo, po, lo, go


def friction_table(schedule_id, L, eps, gamma, steepness):
    s = np.zeros(L)
    g = np.zeros(L)
    lib().oracle_friction_table(int(schedule_id), int(L), float(eps), float(gamma), float(steepness), _ptr(s), _ptr(g))
    return s, g


def welford_finalize(m2, count):
    m2 = np.ascontiguousarray(m2, dtype=np.float64)
    Cn, D = m2.shape
    out = np.zeros(D)
    lib().oracle_welford_finalize(_ptr(m2), float(count), Cn, D, _ptr(out))
    return out


def philox(ctr, key):
    c = np.asarray(ctr, dtype=np.uint32)
    k = np.asarray(key, dtype=np.uint32)
    o = np.zeros(4, dtype=np.uint32)
    lib().oracle_philox4x32_10(_ptr(c), _ptr(k), _ptr(o))
    return o
