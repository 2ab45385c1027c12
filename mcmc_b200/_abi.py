"""ctypes mirror of include/mcmc_b200.h and the loader of libmcmc_b200.so.

There is no CPU fallback: if the shared library is missing or cannot be
loaded, every entry raises (the product path must fail loudly).
"""
from __future__ import annotations

import ctypes as C
import os

ABI_VERSION = 2

SAMPLER_HMC, SAMPLER_GRAHMC, SAMPLER_RWMH, SAMPLER_NUTS = 0, 1, 2, 3

TARGET_IDS = {
    "standard_normal": 0,
    "correlated_gaussian": 1,
    "ill_conditioned_gaussian": 2,
    "neals_funnel": 3,
    "log_gamma": 4,
    "student_t": 5,
    "rosenbrock": 6,
    "gaussian_mixture": 7,
    "dense_gaussian": 8,
}
SCHEDULE_IDS = {"constant": 0, "tanh": 1, "sigmoid": 2, "linear": 3, "sine": 4}

RNG_PHILOX, RNG_INJECTED = 0, 1
FLAG_TRACK_PROPOSALS, FLAG_ACCUM_WELFORD, FLAG_NO_SAMPLES = 1, 2, 4
MAX_DIM = 128
MAX_TREE_DEPTH = 20

OK, EINVAL, ECUDA, ENOTSUP, EWORKSPACE = 0, -1, -2, -3, -4


class McmcParams(C.Structure):
    _fields_ = [
        ("abi_version", C.c_uint32),
        ("target_id", C.c_int32),
        ("schedule_id", C.c_int32),
        ("rng_mode", C.c_int32),
        ("target_params", C.c_double * 8),
        ("n_chains", C.c_int64),
        ("num_samples", C.c_int64),
        ("burn_in", C.c_int64),
        ("chain_offset", C.c_int64),
        ("n_dim", C.c_int32),
        ("num_steps", C.c_int32),
        ("max_tree_depth", C.c_int32),
        ("flags", C.c_uint32),
        ("step_size", C.c_double),
        ("gamma", C.c_double),
        ("steepness", C.c_double),
        ("delta_max", C.c_double),
        ("seed", C.c_uint32 * 2),
        ("reserved", C.c_uint32 * 2),
    ]


_P = C.c_void_p


class McmcBuffers(C.Structure):
    _fields_ = [
        ("init_q", _P), ("inv_mass", _P), ("cov_inv", _P), ("z_inject", _P), ("u_inject", _P),
        ("nuts_dir_u", _P), ("nuts_prop_u", _P),
        ("samples", _P), ("log_probs", _P), ("accept_count", _P), ("accept_rate", _P),
        ("final_q", _P), ("final_lp", _P), ("final_g", _P),
        ("pre_q", _P), ("pre_lp", _P), ("prop_q", _P), ("prop_lp", _P), ("delta_h", _P),
        ("tree_depths", _P), ("mean_accept_probs", _P),
        ("welford_mean", _P), ("welford_m2", _P), ("welford_count", C.c_double),
        ("counters", _P), ("workspace", _P), ("workspace_bytes", C.c_size_t),
        ("step_size_dev", _P),
    ]


BUFFER_FIELDS = [f[0] for f in McmcBuffers._fields_ if f[1] is _P and f[0] != "workspace"]

# every symbol include/mcmc_b200.h declares
EXPORTS = [
    "mcmc_abi_version", "mcmc_build_info", "mcmc_last_error", "mcmc_device_count", "mcmc_set_device", "mcmc_workspace_bytes",
    "mcmc_hmc_run", "mcmc_grahmc_run", "mcmc_rwmh_run", "mcmc_nuts_run",
    "mcmc_run_host", "mcmc_host_arena_release", "mcmc_register_user_target", "mcmc_target_eval", "mcmc_leapfrog", "mcmc_philox_fill", "mcmc_friction_table",
    "mcmc_welford_finalize", "mcmc_welford_finalize_noalloc", "mcmc_chain_moments", "mcmc_da_init", "mcmc_da_update", "mcmc_da_reset", "mcmc_sum_f64", "mcmc_fp64_peak_probe",
    "mcmc_rank_normalize", "mcmc_rank_normalize_fold_workspace", "mcmc_rank_normalize_fold", "mcmc_split_chain_stats_workspace", "mcmc_split_chain_stats",
]

_LIBS: dict = {}
_CSRC = os.path.join(os.path.dirname(os.path.abspath(__file__)), "csrc")
LIB_PATHS = {
    "fast": os.path.join(_CSRC, "libmcmc_b200.so"),
    # same sources, -DMCMC_FAITHFUL -fmad=false: the reference's operation order, one rounding per op
    "strict": os.path.join(_CSRC, "libmcmc_b200_strict.so"),
}
LIB_PATH = LIB_PATHS["fast"]


class McmcError(RuntimeError):
    def __init__(self, code: int, msg: str):
        super().__init__(f"mcmc_b200 error {code}: {msg}")
        self.code = code


def default_variant() -> str:
    return "strict" if os.environ.get("MCMC_B200_STRICT", "0") not in ("", "0") else "fast"


def lib(variant: str | None = None) -> C.CDLL:
    """Load libmcmc_b200[_strict].so (built by __graft_entry__.build()); raise if absent."""
    variant = variant or default_variant()
    if variant in _LIBS:
        return _LIBS[variant]
    path = LIB_PATHS[variant]
    if variant == "fast" and os.environ.get("MCMC_B200_LIB"):   # kernel experiments: an alternative build of the same ABI
        path = os.environ["MCMC_B200_LIB"]
    if not os.path.exists(path):
        raise ImportError(
            f"{path} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
            "(there is no CPU fallback)")
    L = C.CDLL(path)
    L.mcmc_abi_version.restype = C.c_int
    L.mcmc_last_error.restype = C.c_char_p
    L.mcmc_build_info.restype = C.c_char_p
    L.mcmc_device_count.restype = C.c_int
    L.mcmc_set_device.argtypes = [C.c_int]
    L.mcmc_workspace_bytes.argtypes = [C.c_int, C.POINTER(McmcParams), C.POINTER(C.c_size_t)]
    for name in ("mcmc_hmc_run", "mcmc_grahmc_run", "mcmc_rwmh_run", "mcmc_nuts_run"):
        getattr(L, name).argtypes = [C.POINTER(McmcParams), C.POINTER(McmcBuffers), _P]
    L.mcmc_run_host.argtypes = [C.c_int, C.POINTER(McmcParams), C.POINTER(McmcBuffers), C.c_int]
    L.mcmc_target_eval.argtypes = [C.POINTER(McmcParams), _P, _P, C.c_int64, _P, _P, _P]
    L.mcmc_register_user_target.argtypes = [_P, _P, _P, _P, _P]
    L.mcmc_leapfrog.argtypes = [C.POINTER(McmcParams), C.c_int] + [_P] * 11
    L.mcmc_philox_fill.argtypes = [C.POINTER(McmcParams), C.c_int64, C.c_int64, _P, _P, _P]
    L.mcmc_friction_table.argtypes = [C.c_int, C.c_int, C.c_double, C.c_double, C.c_double, _P, _P, _P]
    L.mcmc_welford_finalize.argtypes = [_P, C.c_double, C.c_int64, C.c_int32, _P, _P, C.c_size_t, _P]
    L.mcmc_welford_finalize_noalloc.argtypes = [_P, C.c_double, C.c_int64, C.c_int32, _P, _P]
    L.mcmc_chain_moments.argtypes = [_P, C.c_int64, C.c_int64, C.c_int32, _P, _P, C.c_double, _P]
    L.mcmc_da_init.argtypes = [_P, C.c_double, _P]
    L.mcmc_da_update.argtypes = [_P, _P, C.c_double, C.c_double, C.c_double, C.c_double, C.c_double, _P]
    L.mcmc_da_reset.argtypes = [_P, _P]
    L.mcmc_sum_f64.argtypes = [_P, C.c_int64, _P, _P, C.c_size_t, _P]
    L.mcmc_fp64_peak_probe.argtypes = [C.c_int, C.POINTER(C.c_double), C.POINTER(C.c_float), _P]
    L.mcmc_rank_normalize.argtypes = [_P, _P, C.c_int64, _P, _P]
    L.mcmc_rank_normalize_fold_workspace.argtypes = [C.c_int64, C.POINTER(C.c_size_t)]
    L.mcmc_rank_normalize_fold.argtypes = [_P, C.c_int64, _P, _P, _P, _P, C.c_size_t, _P]
    L.mcmc_split_chain_stats_workspace.argtypes = [C.c_int64, C.c_int64, C.c_int32, C.POINTER(C.c_size_t)]
    L.mcmc_split_chain_stats.argtypes = [_P, C.c_int64, C.c_int64, C.c_int64, C.c_int64, C.c_int64, C.c_int32, _P, _P,
                                         _P, C.c_size_t, _P]
    if L.mcmc_abi_version() != ABI_VERSION:
        raise ImportError(f"ABI mismatch: library {L.mcmc_abi_version()} vs python {ABI_VERSION}")
    _LIBS[variant] = L
    return L


def check(rc: int, variant: str | None = None) -> None:
    """Map a non-zero return code to the exception the reference would raise."""
    if rc != 0:
        msg = lib(variant).mcmc_last_error()
        text = msg.decode() if msg else ""
        if rc == EINVAL:
            raise ValueError(text)
        raise McmcError(rc, text)
