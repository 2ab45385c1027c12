"""Loader for tests/golden/reference_leapfrog_golden.npz — the reference's own `leapfrog` (samplers/HMC.py:85-134) and
`_trajectory_with_schedule` (samplers/GRAHMC.py:215-261) run by tests/golden/make_reference_leapfrog_golden.py — and the
mapping of a case onto the C ABI's parameters (shared by the oracle test and the CUDA test)."""
import ast
import os

import numpy as np

from mcmc_b200 import _abi
from mcmc_b200.benchmarks import targets as T
from reference_cases import _TARGET_KW

PATH = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "reference_leapfrog_golden.npz")
OUT_KEYS = ("q1", "p1", "lp1", "g1")


def load():
    blob = np.load(PATH, allow_pickle=False)
    cases = {}
    for cid in blob["ids"]:
        cid = str(cid)
        c = {k.split("|", 1)[1]: blob[k] for k in blob.files if k.startswith(cid + "|")}
        c["meta"] = ast.literal_eval(str(c["meta"]))
        cases[cid] = c
    return cases


def params_of(c, make_params):
    """(McmcParams, tag, friction flag) of a fixture case."""
    m = c["meta"]
    tag = T.resolve_log_prob(getattr(T, m["target"])(dim=m["D"], **_TARGET_KW[m["target"]]).log_prob_fn)
    friction = m["fn"] == "trajectory"
    p = make_params(tag.target_id, tag.target_params, m["C"], m["D"], 0, 0, m["L"], float(c["eps"]),
                    gamma=m.get("gamma", 0.0), steepness=m.get("steep", 1.0),
                    schedule_id=_abi.SCHEDULE_IDS[m.get("sched", "constant")])
    return p, tag, friction


def run_oracle(O, c):
    p, tag, friction = params_of(c, O.make_params)
    q, mom, lp, g = O.leapfrog(p, c["q0"], c["p0"], c["lp0"], c["g0"], inv_mass=c["inv_mass"], cov_inv=tag.cov_inv, friction=friction)
    return dict(q1=q, p1=mom, lp1=lp, g1=g)


def run_cuda(c, variant):
    """Through the reference-named Python functions (samplers.HMC.leapfrog / samplers.GRAHMC._trajectory_with_schedule),
    which call mcmc_leapfrog of the C ABI."""
    from mcmc_b200.samplers import GRAHMC, HMC
    m = c["meta"]
    f = getattr(T, m["target"])(dim=m["D"], **_TARGET_KW[m["target"]]).log_prob_fn
    if m["fn"] == "leapfrog":
        q, mom, g, lp = HMC.leapfrog(c["q0"], c["p0"], float(c["eps"]), c["lp0"], c["g0"], f, m["L"], c["inv_mass"], _variant=variant)
    else:
        q, mom, lp, g = GRAHMC._trajectory_with_schedule(c["q0"], c["p0"], float(c["eps"]), m["gamma"], m["steep"], c["lp0"], c["g0"],
                                                         m["L"], f, GRAHMC.get_friction_schedule(m["sched"]), c["inv_mass"],
                                                         _variant=variant)
    return dict(q1=q.cpu().numpy(), p1=mom.cpu().numpy(), lp1=lp.cpu().numpy(), g1=g.cpu().numpy())
