"""Generates tests/golden/reference_golden.npz: outputs of the REFERENCE'S OWN SOURCE on small seeded cases.

jax 0.8.0 cannot be installed here (no wheel, no network), so the reference's files are imported unmodified from
/root/reference (samplers/HMC.py, GRAHMC.py, RWMH.py, NUTS.py, benchmarks/targets.py) on top of the torch-fp64
stand-in for the jax API in tests/golden/jax_shim (eager IEEE double arithmetic in the order the reference's source
states it; vmap = loop over chains, value_and_grad = torch.autograd, lax.* = Python control flow).  What this pins:
every formula, operation order, state update, burn-in / counter reset, output stacking and the NUTS tree logic of the
reference as WRITTEN.  What it does not pin: XLA's own exp/log/tanh roundings and jax's threefry streams — the draws
are the stand-in's and are recorded here so the oracle and the CUDA kernels receive exactly the same momenta and
uniforms (the north star's "reference's own momenta and uniforms injected" check).

Run in the build container only (/root/reference does not exist on the GPU box):
    python tests/golden/make_reference_golden.py
"""
import importlib.util
import os
import sys
import types

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REF = os.environ.get("MCMC_REFERENCE", "/root/reference")
sys.path.insert(0, os.path.join(HERE, "jax_shim"))
for name in ("matplotlib", "matplotlib.pyplot"):   # GRAHMC.py imports pyplot at module level (plots are not on the path)
    sys.modules.setdefault(name, types.ModuleType(name))

import jax  # noqa: E402  (the stand-in)
import jax.numpy as jnp  # noqa: E402
from jax import random  # noqa: E402
from jax._core import DRAW_LOG  # noqa: E402
import torch  # noqa: E402


def load(relpath, name):
    spec = importlib.util.spec_from_file_location(name, os.path.join(REF, relpath))
    mod = importlib.util.module_from_spec(spec)
    sys.modules[name] = mod
    spec.loader.exec_module(mod)
    return mod


T = load("benchmarks/targets.py", "ref_targets")
HMC = load("samplers/HMC.py", "ref_HMC")
GRAHMC = load("samplers/GRAHMC.py", "ref_GRAHMC")
RWMH = load("samplers/RWMH.py", "ref_RWMH")
NUTS = load("samplers/NUTS.py", "ref_NUTS")

# (target factory name, kwargs, step size, init) — same conditioning as tests/helpers.TARGET_CASES
TARGETS = {
    "standard_normal": ({}, 0.2, lambda r, C, D: r.standard_normal((C, D))),
    "correlated_gaussian": (dict(correlation=0.9), 0.05, lambda r, C, D: r.standard_normal((C, D))),
    "ill_conditioned_gaussian": (dict(condition_number=100.0), 0.3, lambda r, C, D: r.standard_normal((C, D))),
    "neals_funnel": ({}, 0.05, lambda r, C, D: r.standard_normal((C, D))),
    "log_gamma": (dict(shape=2.0, rate=1.0), 0.05, lambda r, C, D: 0.5 + np.abs(r.standard_normal((C, D)))),
    "student_t": (dict(df=3.0), 0.2, lambda r, C, D: 2.0 * r.standard_normal((C, D))),
    "rosenbrock": (dict(scale=0.1), 0.005, lambda r, C, D: 1.0 + 0.05 * r.standard_normal((C, D))),
    "gaussian_mixture": (dict(n_modes=2, separation=5.0), 0.2, lambda r, C, D: r.standard_normal((C, D))),
}

CASES = []
for _n in TARGETS:
    CASES.append(dict(id=f"hmc/{_n}", sampler="hmc", target=_n, C=5, D=20, S=5, B=2, L=8, seed=11))
CASES.append(dict(id="hmc/rosenbrock/L32_unit_mass", sampler="hmc", target="rosenbrock", C=3, D=20, S=3, B=0, L=32, seed=12,
                  unit_mass=True))
CASES.append(dict(id="hmc/standard_normal/D10_1d_init", sampler="hmc", target="standard_normal", C=1, D=10, S=4, B=1, L=4,
                  seed=13, one_d=True))
for _s, _st in (("constant", 2.0), ("tanh", 0.5), ("sigmoid", 2.0), ("linear", 2.0), ("sine", 2.0)):
    CASES.append(dict(id=f"grahmc/{_s}/neals_funnel", sampler="grahmc", target="neals_funnel", sched=_s, steep=_st, gamma=1.0,
                      C=5, D=20, S=5, B=2, L=8, seed=21))
CASES.append(dict(id="grahmc/tanh/rosenbrock", sampler="grahmc", target="rosenbrock", sched="tanh", steep=0.5, gamma=0.7,
                  C=4, D=20, S=4, B=1, L=16, seed=22))
CASES.append(dict(id="grahmc/none/standard_normal_D10", sampler="grahmc", target="standard_normal", sched=None, steep=None,
                  gamma=1.0, C=4, D=10, S=6, B=3, L=16, seed=23, eps=0.05))
for _n in ("standard_normal", "gaussian_mixture", "student_t"):
    CASES.append(dict(id=f"rwmh/{_n}", sampler="rwmh", target=_n, C=6, D=20, S=8, B=3, L=0, seed=31))
for _n, _md in (("standard_normal", 6), ("rosenbrock", 5), ("neals_funnel", 5), ("correlated_gaussian", 5)):
    CASES.append(dict(id=f"nuts/{_n}", sampler="nuts", target=_n, C=3, D=20, S=4, B=1, L=0, md=_md, seed=41))
# edge cases: trajectories that overflow (non-finite guard HMC.py:199, reject), NUTS divergence (delta_max, NUTS.py:283) and the
# depth cap; "state_only": proposals may be inf/nan in implementation-defined ways, so only the chain state is compared
CASES.append(dict(id="hmc/rosenbrock/overflow", sampler="hmc", target="rosenbrock", C=6, D=20, S=4, B=1, L=8, seed=51, eps=0.45,
                  state_only=True))
CASES.append(dict(id="grahmc/linear/rosenbrock/overflow", sampler="grahmc", target="rosenbrock", sched="linear", steep=2.0, gamma=2.0,
                  C=6, D=20, S=4, B=1, L=8, seed=52, eps=0.45, state_only=True))
CASES.append(dict(id="nuts/neals_funnel/divergent", sampler="nuts", target="neals_funnel", C=4, D=20, S=4, B=1, L=0, md=6, seed=53,
                  eps=2.5, state_only=True))
CASES.append(dict(id="nuts/standard_normal/depth_cap2", sampler="nuts", target="standard_normal", C=4, D=20, S=5, B=0, L=0, md=2, seed=54,
                  eps=0.05))


def np64(x):
    return np.ascontiguousarray(torch.as_tensor(x).detach().to(torch.float64).numpy())


def run_case(c):
    kw, eps, init = TARGETS[c["target"]]
    eps = c.get("eps", eps)
    if c["sampler"] == "nuts" and "eps" not in c:
        eps = {"standard_normal": 0.4, "rosenbrock": 0.01, "neals_funnel": 0.1, "correlated_gaussian": 0.1}[c["target"]]
    if c["sampler"] == "rwmh":
        eps = 2.38 / np.sqrt(c["D"])   # dual_averaging.py:53
    rng = np.random.default_rng(c["seed"])
    C_, D, S, B, L = c["C"], c["D"], c["S"], c["B"], c["L"]
    q0 = init(rng, C_, D)
    w = np.ones(D) if c.get("unit_mass") else np.linspace(0.5, 2.0, D)
    tgt = getattr(T, c["target"])(dim=D, **kw)
    f = tgt.log_prob_fn
    key = random.PRNGKey(c["seed"])
    q0j = jnp.asarray(q0[0] if c.get("one_d") else q0)
    wj = None if c.get("unit_mass") else jnp.asarray(w)
    del DRAW_LOG[:]
    out = {}
    if c["sampler"] == "hmc":
        r = HMC.hmc_run(key, f, q0j, eps, L, S, burn_in=B, inv_mass_matrix=wj, track_proposals=True)
    elif c["sampler"] == "grahmc":
        sched = None if c["sched"] is None else GRAHMC.get_friction_schedule(c["sched"])
        r = GRAHMC.rahmc_run(key, f, q0j, eps, L, c["gamma"], c["steep"], S, burn_in=B, inv_mass_matrix=wj,
                             friction_schedule=sched, track_proposals=True)
    elif c["sampler"] == "rwmh":
        r = RWMH.rwMH_run(key, f, q0j, S, eps, burn_in=B)
    else:
        r = NUTS.nuts_run(key, f, q0j, eps, S, burn_in=B, inv_mass_matrix=wj, max_tree_depth=c["md"], delta_max=1000.0)
    log = list(DRAW_LOG)
    out["samples"], out["log_probs"] = np64(r[0]), np64(r[1])
    out["accept_rate"] = np64(r[2]).astype(np.float32)
    st = r[3]
    out["final_q"], out["final_lp"] = np64(st.position), np64(st.log_prob)
    if hasattr(st, "grad_log_prob"):
        out["final_g"] = np64(st.grad_log_prob)
    out["accept_count"] = np64(st.accept_count).astype(np.int32)
    T_ = B + S
    if c["sampler"] in ("hmc", "grahmc"):
        out["pre_q"], out["pre_lp"], out["prop_q"], out["prop_lp"], out["delta_h"] = (np64(x) for x in r[4:9])
    if c["sampler"] == "nuts":
        out["tree_depths"] = np64(r[4]).astype(np.int32)
        out["mean_accept_probs"] = np64(r[5])
    # ---- the draws, in the layout the C ABI's injected mode takes
    ins = dict(q0=q0, inv_mass=w, eps=np.float64(eps))
    if c["sampler"] != "nuts":
        assert len(log) == 2 * T_ and all(k == ("normal", "uniform")[i % 2] for i, (k, _, _) in enumerate(log)), c["id"]
        ins["z"] = np.stack([log[2 * i][2].reshape(C_, D) for i in range(T_)])
        ins["u"] = np.stack([log[2 * i + 1][2].reshape(C_) for i in range(T_)])
    else:
        md = c["md"]
        z = np.zeros((T_, C_, D)); u = np.zeros((T_, C_)); du = np.full((T_, C_, md), 0.25); pu = np.full((T_, C_, md), 0.75)
        # per iteration, per chain (vmap = loop over chains): normal (D,), slice uniform, then one (bernoulli, uniform) pair per doubling
        pos = 0
        for t in range(T_):
            for ch in range(C_):
                assert log[pos][0] == "normal" and log[pos + 1][0] == "uniform", (c["id"], pos)
                z[t, ch] = log[pos][2]; u[t, ch] = log[pos + 1][2]; pos += 2
                d = 0
                while pos < len(log) and log[pos][0] == "bernoulli":
                    assert log[pos + 1][0] == "uniform"
                    du[t, ch, d] = log[pos][2]; pu[t, ch, d] = log[pos + 1][2]; pos += 2; d += 1
                if t >= B:
                    assert d == out["tree_depths"][t - B, ch], (c["id"], t, ch, d)
        assert pos == len(log)
        ins.update(z=z, u=u, dir_u=du, prop_u=pu)
    return ins, out


if __name__ == "__main__":
    blob = {}
    ids = []
    for c in CASES:
        ins, out = run_case(c)
        ids.append(c["id"])
        for k, v in {**ins, **out}.items():
            blob[f"{c['id']}|{k}"] = v
        meta = {k: v for k, v in c.items() if k != "id"}
        blob[f"{c['id']}|meta"] = np.array(repr(meta))
        print(f"{c['id']:44s} accept {out['accept_rate'].mean():.3f}  lp[-1] {out['log_probs'][-1].mean():+.4f}"
              + (f"  depths {out['tree_depths'].min()}..{out['tree_depths'].max()}" if "tree_depths" in out else ""))
    blob["ids"] = np.array(ids)
    path = os.path.join(HERE, "reference_golden.npz")
    np.savez_compressed(path, **blob)
    print("wrote", path, os.path.getsize(path), "bytes")
