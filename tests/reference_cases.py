"""Loader for tests/golden/reference_golden.npz — outputs of the reference's own source files (run on the torch-fp64
stand-in for jax by tests/golden/make_reference_golden.py) — and the mapping of a case onto the C ABI's parameters."""
import ast
import os

import numpy as np

from mcmc_b200 import _abi
from mcmc_b200.benchmarks import targets as T

PATH = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "reference_golden.npz")
_TARGET_KW = {
    "standard_normal": {}, "correlated_gaussian": dict(correlation=0.9), "ill_conditioned_gaussian": dict(condition_number=100.0),
    "neals_funnel": {}, "log_gamma": dict(shape=2.0, rate=1.0), "student_t": dict(df=3.0), "rosenbrock": dict(scale=0.1),
    "gaussian_mixture": dict(n_modes=2, separation=5.0),
}
SAMPLERS = {"hmc": _abi.SAMPLER_HMC, "grahmc": _abi.SAMPLER_GRAHMC, "rwmh": _abi.SAMPLER_RWMH, "nuts": _abi.SAMPLER_NUTS}
FLOAT_KEYS = ("samples", "log_probs", "final_q", "final_lp", "final_g", "pre_q", "pre_lp", "prop_q", "prop_lp", "delta_h",
              "mean_accept_probs")
EXACT_KEYS = ("accept_count", "tree_depths")


def load():
    blob = np.load(PATH, allow_pickle=False)
    cases = {}
    for cid in blob["ids"]:
        cid = str(cid)
        c = {k.split("|", 1)[1]: blob[k] for k in blob.files if k.startswith(cid + "|")}
        c["meta"] = ast.literal_eval(str(c["meta"]))
        cases[cid] = c
    return cases


def run_case(backend, c, variant_kw=None):
    """Runs one fixture case through `backend` (the oracle driver, or tests.helpers.GpuBackend) with the recorded draws."""
    m = c["meta"]
    tag = T.resolve_log_prob(getattr(T, m["target"])(dim=m["D"], **_TARGET_KW[m["target"]]).log_prob_fn)
    sampler = SAMPLERS[m["sampler"]]
    flags = _abi.FLAG_TRACK_PROPOSALS if m["sampler"] in ("hmc", "grahmc") else 0
    sched = m.get("sched")
    steep = m.get("steep")
    p = backend.make_params(tag.target_id, tag.target_params, m["C"], m["D"], m["S"], m["B"], m["L"], float(c["eps"]),
                            gamma=m.get("gamma", 0.0), steepness=1.0 if steep is None else steep,       # GRAHMC.py:409
                            schedule_id=_abi.SCHEDULE_IDS["constant" if sched is None else sched],       # GRAHMC.py:395-396
                            rng_mode=_abi.RNG_INJECTED, flags=flags, max_tree_depth=m.get("md", 10), delta_max=1000.0)
    kw = dict(inv_mass=c["inv_mass"], cov_inv=tag.cov_inv, z=c["z"], u=c["u"])
    if m["sampler"] == "nuts":
        kw.update(dir_u=c["dir_u"], prop_u=c["prop_u"])
    return backend.run(sampler, p, c["q0"], **kw)


QUICKSTART_PATH = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "reference_quickstart_golden.npz")


def run_quickstart(backend):
    """BASELINE configs[0] (README.md:41-71) through `backend` with the reference run's recorded draws; returns
    (outputs, fixture)."""
    c = dict(np.load(QUICKSTART_PATH))
    tag = T.resolve_log_prob(T.standard_normal(dim=10, with_norm_const=False).log_prob_fn)   # README's -0.5*sum(x**2)
    p = backend.make_params(tag.target_id, tag.target_params, 4, 10, 1000, 500, 16, 0.05, gamma=1.0, steepness=0.5,
                            schedule_id=_abi.SCHEDULE_IDS["tanh"], rng_mode=_abi.RNG_INJECTED)
    return backend.run(_abi.SAMPLER_GRAHMC, p, c["q0"], z=c["z"], u=c["u"]), c


def check_quickstart(got, c, tol):
    from helpers import relerr
    assert np.array_equal(got["accept_count"], c["accept_count"])
    assert np.array_equal(got["accept_rate"], c["accept_rate"])
    worst = 0.0
    for a, b in ((got["log_probs"], c["log_probs"]), (got["samples"][::20], c["samples_every20"]),
                 (got["samples"][-1], c["samples_last"]), (got["final_q"], c["final_q"]), (got["final_lp"], c["final_lp"]),
                 (got["final_g"], c["final_g"]), (got["samples"].mean(axis=(0, 1)), c["sample_mean"]),
                 (got["samples"].var(axis=(0, 1)), c["sample_var"])):
        e = relerr(a, b)
        worst = max(worst, e)
        assert e <= tol, e
    return worst
