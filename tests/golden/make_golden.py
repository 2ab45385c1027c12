"""Generates tests/golden/hot_path_golden.npz.

The reference has no golden vectors and cannot be imported here (jax missing),
so these fixtures are outputs of the CPU oracle (oracle/mcmc_oracle.c) on seeded
inputs — they freeze the oracle's behaviour (any later edit of the oracle that
changes a number is caught on CPU) and give the GPU parity tests a second,
file-based reference.  They are NOT outputs of the reference itself.

Usage:  python tests/golden/make_golden.py
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
sys.path.insert(0, os.path.dirname(HERE))

from helpers import TARGET_CASES, tid  # noqa: E402
from mcmc_b200 import _abi  # noqa: E402

S_HMC = dict(C=8, D=20, S=6, B=2, L=16)

CASES = []
for _name in TARGET_CASES:
    CASES.append(dict(id=f"hmc/{_name}", sampler=_abi.SAMPLER_HMC, target=_name, seed=100, **S_HMC))
for _sched in _abi.SCHEDULE_IDS:
    CASES.append(dict(id=f"grahmc/{_sched}/neals_funnel", sampler=_abi.SAMPLER_GRAHMC, target="neals_funnel",
                      sched=_sched, seed=200, **S_HMC))
CASES.append(dict(id="grahmc/tanh/quickstart", sampler=_abi.SAMPLER_GRAHMC, target="standard_normal", sched="tanh",
                  seed=201, C=4, D=10, S=8, B=4, L=16, quickstart=True))
CASES.append(dict(id="rwmh/standard_normal", sampler=_abi.SAMPLER_RWMH, target="standard_normal", seed=300,
                  C=8, D=20, S=10, B=3, L=0))
for _name in ("standard_normal", "rosenbrock", "neals_funnel"):
    CASES.append(dict(id=f"nuts/{_name}", sampler=_abi.SAMPLER_NUTS, target=_name, seed=400, C=4, D=20, S=4, B=1,
                      L=0, md=6))


def case_inputs(case):
    params, eps, init = TARGET_CASES[case["target"]]
    rng = np.random.default_rng(case["seed"])
    C_, D, S, B = case["C"], case["D"], case["S"], case["B"]
    q0 = init(rng, C_, D)
    z = rng.standard_normal((B + S, C_, D))
    u = rng.random((B + S, C_))
    md = case.get("md", 10)
    du = rng.random((B + S, C_, md))
    pu = rng.random((B + S, C_, md))
    w = np.linspace(0.5, 2.0, D)
    if case.get("quickstart"):  # README.md:51-67: -0.5*sum(x^2), eps 0.05, gamma 1, steepness 0.5
        params, eps, q0, w = [0.0], 0.05, 2.0 * q0, None
    if case["sampler"] == _abi.SAMPLER_RWMH:
        eps = 2.38 / np.sqrt(D)
        w = None
    return dict(params=params, eps=eps, q0=q0, z=z, u=u, du=du, pu=pu, w=w, md=md)


def run_case(backend, case, **kw):
    """backend: oracle.oracle (CPU) or a GPU adapter with the same make_params/run surface."""
    i = case_inputs(case)
    sampler = case["sampler"]
    flags = _abi.FLAG_TRACK_PROPOSALS if sampler in (_abi.SAMPLER_HMC, _abi.SAMPLER_GRAHMC) else 0
    p = backend.make_params(tid(case["target"]), i["params"], case["C"], case["D"], case["S"], case["B"], case["L"],
                            i["eps"], gamma=1.0, steepness=0.5 if case.get("sched") == "tanh" else 2.0,
                            schedule_id=_abi.SCHEDULE_IDS.get(case.get("sched"), 0), rng_mode=_abi.RNG_INJECTED,
                            flags=flags, max_tree_depth=i["md"])
    r = backend.run(sampler, p, i["q0"], inv_mass=i["w"], z=i["z"], u=i["u"],
                    dir_u=i["du"] if sampler == _abi.SAMPLER_NUTS else None,
                    prop_u=i["pu"] if sampler == _abi.SAMPLER_NUTS else None, **kw)
    keep = ("samples", "log_probs", "accept_count", "final_q", "final_lp", "final_g", "prop_q", "prop_lp", "delta_h",
            "tree_depths", "mean_accept_probs")
    return {k: np.asarray(v) for k, v in r.items() if k in keep}


if __name__ == "__main__":
    from oracle import oracle as O
    O.build()
    out = {}
    for case in CASES:
        for k, v in run_case(O, case).items():
            out[f"{case['id']}/{k}"] = v
    path = os.path.join(HERE, "hot_path_golden.npz")
    np.savez_compressed(path, **out)
    print(path, len(out), "arrays", os.path.getsize(path), "bytes")
