"""Generates tests/golden/reference_leapfrog_golden.npz: the reference's own `leapfrog` (samplers/HMC.py:85-134) and
`_trajectory_with_schedule` (samplers/GRAHMC.py:215-261) run on small seeded inputs.

Same arrangement as make_reference_golden.py: the reference's files are imported unmodified from /root/reference on top
of the torch-fp64 stand-in for jax in tests/golden/jax_shim (eager IEEE double arithmetic in the order the source states
it; value_and_grad = torch.autograd).  It pins the integrator by itself — formulas, operation order, the position of
the friction scalings, the return order of the two functions — for every benchmark target and every friction schedule.
XLA's own exp/log/tanh roundings are not pinned (the stand-in's are torch's).

Run in the build container only (/root/reference does not exist on the GPU box):
    python tests/golden/make_reference_leapfrog_golden.py
"""
import importlib.util
import os
import sys
import types

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REF = os.environ.get("MCMC_REFERENCE", "/root/reference")
sys.path.insert(0, os.path.join(HERE, "jax_shim"))
for name in ("matplotlib", "matplotlib.pyplot"):
    sys.modules.setdefault(name, types.ModuleType(name))

import jax  # noqa: E402,F401  (the stand-in)
import jax.numpy as jnp  # noqa: E402
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

# (factory kwargs, step size, init) — the conditioning of make_reference_golden.py
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
    CASES.append(dict(id=f"leapfrog/{_n}", fn="leapfrog", target=_n, C=4, D=20, L=12, seed=101))
CASES.append(dict(id="leapfrog/rosenbrock/L64_unit_mass", fn="leapfrog", target="rosenbrock", C=3, D=20, L=64, seed=102,
                  unit_mass=True))
CASES.append(dict(id="leapfrog/standard_normal/D10", fn="leapfrog", target="standard_normal", C=5, D=10, L=7, seed=103))
CASES.append(dict(id="leapfrog/neals_funnel/D7_runtime_dim", fn="leapfrog", target="neals_funnel", C=3, D=7, L=9, seed=104))
CASES.append(dict(id="leapfrog/student_t/L0", fn="leapfrog", target="student_t", C=3, D=20, L=0, seed=105))
CASES.append(dict(id="leapfrog/rosenbrock/L1", fn="leapfrog", target="rosenbrock", C=3, D=20, L=1, seed=106))
for _s, _st in (("constant", 2.0), ("tanh", 0.5), ("sigmoid", 2.0), ("linear", 2.0), ("sine", 2.0)):
    CASES.append(dict(id=f"trajectory/{_s}/neals_funnel", fn="trajectory", target="neals_funnel", sched=_s, steep=_st, gamma=1.0,
                      C=4, D=20, L=12, seed=111))
    CASES.append(dict(id=f"trajectory/{_s}/rosenbrock_L5", fn="trajectory", target="rosenbrock", sched=_s, steep=_st, gamma=0.7,
                      C=3, D=20, L=5, seed=112))   # odd L: the constant schedule flips at 2i >= L
CASES.append(dict(id="trajectory/tanh/standard_normal/D10", fn="trajectory", target="standard_normal", sched="tanh", steep=0.5,
                  gamma=1.0, C=4, D=10, L=16, seed=113, eps=0.05))
CASES.append(dict(id="trajectory/sine/gaussian_mixture/L1", fn="trajectory", target="gaussian_mixture", sched="sine", steep=2.0,
                  gamma=1.5, C=3, D=20, L=1, seed=114))


def np64(x):
    return np.ascontiguousarray(torch.as_tensor(x).detach().to(torch.float64).numpy())


def run_case(c):
    kw, eps, init = TARGETS[c["target"]]
    eps = c.get("eps", eps)
    rng = np.random.default_rng(c["seed"])
    C_, D, L = c["C"], c["D"], c["L"]
    q0 = init(rng, C_, D)
    w = np.ones(D) if c.get("unit_mass") else np.linspace(0.5, 2.0, D)
    p0 = rng.standard_normal((C_, D)) / np.sqrt(w)
    tgt = getattr(T, c["target"])(dim=D, **kw)
    f = tgt.log_prob_fn
    # value and gradient at the start exactly as hmc_init takes them (HMC.py:73)
    st = HMC.hmc_init(jnp.asarray(q0), f)
    lp0, g0 = st.log_prob, st.grad_log_prob
    if c["fn"] == "leapfrog":
        q1, p1, g1, lp1 = HMC.leapfrog(jnp.asarray(q0), jnp.asarray(p0), eps, lp0, g0, f, L, jnp.asarray(w))
    else:
        sched = GRAHMC.get_friction_schedule(c["sched"])
        q1, p1, lp1, g1 = GRAHMC._trajectory_with_schedule(jnp.asarray(q0), jnp.asarray(p0), eps, c["gamma"], c["steep"], lp0, g0,
                                                           L, f, sched, jnp.asarray(w))
    ins = dict(q0=q0, p0=p0, lp0=np64(lp0), g0=np64(g0), inv_mass=w, eps=np.float64(eps))
    out = dict(q1=np64(q1), p1=np64(p1), lp1=np64(lp1), g1=np64(g1))
    return ins, out


if __name__ == "__main__":
    blob = {}
    ids = []
    for c in CASES:
        ins, out = run_case(c)
        ids.append(c["id"])
        for k, v in {**ins, **out}.items():
            blob[f"{c['id']}|{k}"] = v
        blob[f"{c['id']}|meta"] = np.array(repr({k: v for k, v in c.items() if k != "id"}))
        print(f"{c['id']:44s} lp0 {ins['lp0'].mean():+.4f} -> lp1 {out['lp1'].mean():+.4f}   |q1-q0| {np.abs(out['q1'] - ins['q0']).max():.3e}")
    blob["ids"] = np.array(ids)
    path = os.path.join(HERE, "reference_leapfrog_golden.npz")
    np.savez_compressed(path, **blob)
    print("wrote", path, os.path.getsize(path), "bytes")
