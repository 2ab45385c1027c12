"""Generates tests/golden/reference_host_golden.npz: outputs of the reference's own HOST-SIDE source on seeded inputs —
dual averaging (tuning/dual_averaging.py da_* and joint_da_*), Welford (tuning/welford.py), the warmup schedule
(tuning/adaptation.py build_schedule), the five friction schedules (samplers/GRAHMC.py:69-106), and every
benchmarks/targets.py factory: true_mean / true_cov, log-density and autodiff gradient at random points for default and
non-default parameters.  Same mechanism as make_reference_golden.py (the reference's files imported unmodified from
/root/reference on the torch-fp64 stand-in for jax).  Build container only."""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
import make_reference_golden as G  # noqa: E402  (sets up the stand-in and loads samplers / targets)

sys.path.insert(0, G.REF)          # tuning/*.py import `samplers.X` / `tuning.Y` as packages
import jax  # noqa: E402
import jax.numpy as jnp  # noqa: E402
import torch  # noqa: E402
from tuning import adaptation as A, dual_averaging as DA, welford as W  # noqa: E402

np64 = G.np64
blob = {}

# ---- dual averaging: 45 updates with two window resets (adaptation.py:164-165 pattern)
rng = np.random.default_rng(5)
for tag, eps0, target in (("da_hmc", 0.1, 0.65), ("da_nuts", 0.5, 0.8), ("da_rwmh", 2.38 / np.sqrt(20), 0.234)):
    acc = np.clip(target + 0.25 * rng.standard_normal(45), 0.0, 1.0)
    st = DA.da_init(eps0)
    rows = [[st.log_step, st.log_step_bar, st.H_bar, st.mu, st.count]]
    for i, a in enumerate(acc):
        if i in (12, 30):
            st = DA.da_reset(st)
            rows.append([st.log_step, st.log_step_bar, st.H_bar, st.mu, st.count])
        st = DA.da_update(st, float(a), target)
        rows.append([st.log_step, st.log_step_bar, st.H_bar, st.mu, st.count])
    blob[f"{tag}|accept"] = acc
    blob[f"{tag}|init"] = np.array([eps0, target])
    blob[f"{tag}|states"] = np.array(rows, dtype=np.float64)
acc = np.clip(0.65 + 0.3 * rng.standard_normal(40), 0.0, 1.0)
st = DA.joint_da_init(jnp.array([0.1, 1.0]))
rows = []
for a in acc:
    st = DA.joint_da_update(st, float(a), 0.65)
    rows.append(np.concatenate([np64(st.log_params), np64(st.log_params_bar), [st.H_bar, st.count]]))
blob["joint_da|accept"] = acc
blob["joint_da|states"] = np.array(rows)

# ---- Welford
x = rng.standard_normal((60, 7)) * np.linspace(0.1, 5.0, 7) + np.arange(7)
st = W.welford_init(7)
for row in x[:20]:
    st = W.welford_update(st, jnp.asarray(row))
blob["welford|x"] = x
blob["welford|after20"] = np.stack([np64(st.mean), np64(st.m2)])
st = W.welford_update_batch(st, jnp.asarray(x[20:]))
mean, var = W.welford_covariance(st)
blob["welford|final"] = np.stack([np64(st.mean), np64(st.m2), np64(mean), np64(var)])
blob["welford|count"] = np64(st.count)
m1, v1 = W.welford_covariance(W.welford_update(W.welford_init(7), jnp.asarray(x[0])))   # n clamps to 2
blob["welford|one_sample_var"] = np64(v1)

# ---- warmup schedule
for tag, kw in (("default", {}), ("short", dict(exploration_steps=100, adaptation_windows=[25, 50], cooldown_steps=25))):
    s = A.build_schedule(**kw)
    blob[f"schedule|{tag}"] = np.array([[a, b, {"exploration": 0, "adaptation": 1, "cooldown": 2}[p]] for a, b, p in s])

# ---- friction schedules gamma(t) (GRAHMC.py:69-106): t = i*eps, T = eps*L
for name in ("constant", "tanh", "sigmoid", "linear", "sine"):
    f = G.GRAHMC.get_friction_schedule(name)
    for L, eps, gam, steep in ((16, 0.05, 1.0, 0.5), (7, 0.3, 2.5, 2.0), (128, 0.01, 0.3, 5.0)):
        T_ = eps * L
        blob[f"friction|{name}|{L}|{eps}|{gam}|{steep}"] = np.array(
            [float(f(jnp.asarray(i * eps), jnp.asarray(T_), gam, steep)) for i in range(L)])

# ---- targets: metadata, value and autodiff gradient
TARGET_PARAMS = {
    "standard_normal": [{}],
    "correlated_gaussian": [dict(correlation=0.9), dict(correlation=0.3)],
    "ill_conditioned_gaussian": [dict(condition_number=100.0), dict(condition_number=1e4)],
    "neals_funnel": [{}],
    "log_gamma": [dict(shape=2.0, rate=1.0), dict(shape=3.5, rate=0.5)],
    "student_t": [dict(df=3.0), dict(df=7.5)],
    "rosenbrock": [dict(scale=0.1), dict(scale=0.02)],
    "gaussian_mixture": [dict(n_modes=2, separation=5.0), dict(n_modes=2, separation=2.0)],
}
vg = jax.vmap(jax.value_and_grad(lambda x: x.sum()))  # placeholder to warm the stand-in
names = []
for name, plist in TARGET_PARAMS.items():
    for kw in plist:
        for D in (10, 20, 7) if name != "rosenbrock" else (10, 20, 6):
            t = getattr(G.T, name)(dim=D, **kw)
            if name == "log_gamma":
                pts = 0.2 + np.abs(rng.standard_normal((12, D))) * 1.5
            else:
                pts = rng.standard_normal((12, D)) * (2.0 if name == "student_t" else 1.0)
            lp, g = jax.vmap(jax.value_and_grad(t.log_prob_fn))(jnp.asarray(pts))
            key = f"target|{name}|{D}|{repr(sorted(kw.items()))}"
            names.append(key)
            blob[key + "|x"] = pts
            blob[key + "|lp"] = np64(lp)
            blob[key + "|grad"] = np64(g)
            blob[key + "|true_mean"] = np64(t.true_mean) if t.true_mean is not None else np.zeros(0)
            blob[key + "|true_cov"] = np64(t.true_cov) if t.true_cov is not None else np.zeros(0)
blob["target_keys"] = np.array(names)
# known-answer facts of SURVEY.md 8c, from the reference's own functions
blob["kat|rosenbrock_at_ones"] = np64(G.T.rosenbrock(dim=20).log_prob_fn(jnp.ones(20)))
blob["kat|standard_normal_at_zero"] = np64(G.T.standard_normal(dim=20).log_prob_fn(jnp.zeros(20)))

# ---- Sliced Wasserstein-2 (benchmarks/metrics.py:12-84) with the projection directions recorded
from benchmarks import metrics as M  # noqa: E402
from jax._core import DRAW_LOG  # noqa: E402
s1 = rng.standard_normal((300, 5)) * np.array([1.0, 2.0, 0.5, 1.0, 3.0])
s2 = rng.standard_normal((300, 5)) + 0.3
s3 = rng.standard_normal((200, 5)) * 1.5
for tag, a, b in (("equal", s1, s2), ("unequal", s1, s3)):
    del DRAW_LOG[:]
    val = M.sliced_wasserstein_distance(jnp.asarray(a), jnp.asarray(b), n_projections=30, projection_batch_size=16)
    blob[f"sw2|{tag}|a"], blob[f"sw2|{tag}|b"] = a, b
    blob[f"sw2|{tag}|directions"] = np.concatenate([v.reshape(-1, 5) for k, _, v in DRAW_LOG if k == "normal"])
    blob[f"sw2|{tag}|value"] = np.float64(val)

# ---- Bonferroni z-test gate and gate constants (run_benchmarks.py:55-62,90-158); arviz is only imported there, stubbed
import types  # noqa: E402
import pandas as pd  # noqa: E402
sys.modules.setdefault("arviz", types.ModuleType("arviz"))
RB = G.load("run_benchmarks.py", "ref_run_benchmarks")
for name in ("MIN_ESS_HARD_GATE", "MIN_ESS_TAIL_HARD_GATE", "MIN_ESS_QUALITY", "MIN_ESS_TAIL_QUALITY", "INEFFICIENT_THRESHOLD",
             "HIGH_EFFICIENCY_THRESHOLD"):
    blob[f"gate_const|{name}"] = np.float64(getattr(RB, name))
for tag, D, shift in (("pass", 20, 1.0), ("fail", 20, 4.5), ("d10", 10, 2.9)):
    tgt = G.T.ill_conditioned_gaussian(dim=D)
    mcse = 0.01 + 0.02 * rng.random(D)
    means = np64(tgt.true_mean) + shift * mcse * np.sign(rng.standard_normal(D)) * rng.random(D)
    r = RB.check_summary_statistics({"summary": pd.DataFrame({"mean": means, "mcse_mean": mcse})}, tgt)
    blob[f"ztest|{tag}|means"], blob[f"ztest|{tag}|mcse"] = means, mcse
    blob[f"ztest|{tag}|out"] = np.array([float(r["pass"]), r["max_z"], r["threshold"]])
    blob[f"ztest|{tag}|z"] = np.asarray(r["z_scores"], dtype=np.float64)
dh = np.array([0.0, 999.9, -1000.1, 1e4, -3.0, np.inf])
blob["divergences|dh"] = dh
blob["divergences|flags"] = np64(RB.detect_divergences(jnp.asarray(dh))).astype(np.int8)

# ---- run_adaptive_warmup control flow (tuning/adaptation.py:72-348) with the sampler replaced by warmup_mock.mock_batch
import collections  # noqa: E402
import contextlib  # noqa: E402
import io  # noqa: E402
from warmup_mock import mock_batch  # noqa: E402
_State = collections.namedtuple("_State", "position")
for tag, sampler, kw in (("hmc", "hmc", {}), ("nuts", "nuts", {}), ("hmc_identity", "hmc", dict(learn_mass_matrix=False))):
    calls = [0]
    steps = []

    def fake_run(key, f, position, step_size, *a, num_samples=None, inv_mass_matrix=None, **k):
        smp, cnt = mock_batch(calls[0], np64(position), step_size, None if inv_mass_matrix is None else np64(inv_mass_matrix),
                              num_samples)
        calls[0] += 1
        steps.append(float(step_size))
        rate = jnp.asarray(cnt).astype(jnp.float32) / num_samples      # HMC.py:313: float32 count / num_samples
        st = _State(position=jnp.asarray(smp[-1]))
        if sampler == "nuts":                                            # (samples, lps, rate, state, depths, mean_accept_probs)
            return jnp.asarray(smp), None, rate, st, None, jnp.asarray(np.broadcast_to(np64(rate.to(jnp.float64)), smp.shape[:2]).copy())
        return jnp.asarray(smp), None, rate, st

    A.hmc_run, A.nuts_run = fake_run, fake_run
    C_, D_ = 4, 6
    q0 = rng.standard_normal((C_, D_))
    with contextlib.redirect_stdout(io.StringIO()):
        eps, w, pos, info = A.run_adaptive_warmup(sampler, None, None, jnp.asarray(q0), jax.random.PRNGKey(1), num_warmup=2500,
                                                  target_accept=0.65, **kw)
    blob[f"warmup|{tag}|q0"] = q0
    blob[f"warmup|{tag}|step_sizes"] = np.array(steps)
    blob[f"warmup|{tag}|final_step"] = np.float64(eps)
    blob[f"warmup|{tag}|inv_mass"] = np64(w) if w is not None else np.zeros(0)
    blob[f"warmup|{tag}|position"] = np64(pos)

# ---- the convergence loops of dual_averaging_tune_{rwmh,hmc,nuts} (dual_averaging.py:16-382), same mock sampler
for tag in ("rwmh", "hmc", "nuts"):
    calls = [0]

    def fake(key, f, position, *a, num_samples=None, step_size=None, scale=None, **k):
        v = scale if tag == "rwmh" else step_size
        smp, cnt = mock_batch(calls[0], np64(position), v * (0.25 if tag == "rwmh" else 1.0), None, num_samples)
        calls[0] += 1
        rate = jnp.asarray(cnt).astype(jnp.float32) / num_samples
        st = _State(position=jnp.asarray(smp[-1]))
        if tag == "nuts":
            depths = jnp.asarray(np.full(smp.shape[:2], 3 + calls[0] % 2, dtype=np.int32))
            return None, None, rate, st, depths, jnp.asarray(np.broadcast_to(np64(rate.to(jnp.float64)), smp.shape[:2]).copy())
        return None, None, rate, st

    DA.rwMH_run = DA.hmc_run = DA.nuts_run = fake
    q0 = rng.standard_normal((4, 6))
    with contextlib.redirect_stdout(io.StringIO()):
        if tag == "rwmh":
            v, h = DA.dual_averaging_tune_rwmh(jax.random.PRNGKey(2), None, jnp.asarray(q0), max_iter=400)
        elif tag == "hmc":
            v, h = DA.dual_averaging_tune_hmc(jax.random.PRNGKey(2), None, jnp.asarray(q0), 8, max_iter=400)
        else:
            v, h = DA.dual_averaging_tune_nuts(jax.random.PRNGKey(2), None, jnp.asarray(q0), max_iter=400)
    blob[f"tuner|{tag}|q0"] = q0
    blob[f"tuner|{tag}|value"] = np.float64(v)
    blob[f"tuner|{tag}|history"] = np.array(h["scale_history" if tag == "rwmh" else "step_size_history"], dtype=np.float64)
    blob[f"tuner|{tag}|accept"] = np.array(h["accept_history"], dtype=np.float64)
    blob[f"tuner|{tag}|converged_iter"] = np.int64(h["converged_iter"])
    if tag == "nuts":
        blob[f"tuner|{tag}|depth"] = np.array(h["tree_depth_history"], dtype=np.float64)

# ---- joint (step size, gamma) tuning loop (dual_averaging.py:668-804); the mock's accept rate falls with step * (1 + gamma/4)
calls = [0]


def fake_rahmc(key, f, position, *a, num_samples=None, step_size=None, gamma=None, **k):
    smp, cnt = mock_batch(calls[0], np64(position), step_size * (1.0 + 0.25 * gamma), None, num_samples)
    calls[0] += 1
    return None, None, jnp.asarray(cnt).astype(jnp.float32) / num_samples, _State(position=jnp.asarray(smp[-1]))


DA.rahmc_run = fake_rahmc
q0 = rng.standard_normal((4, 6))
with contextlib.redirect_stdout(io.StringIO()):
    st_, gm_, sp_, h = DA.joint_tune_grahmc(jax.random.PRNGKey(3), None, None, jnp.asarray(q0), 8, schedule_type="tanh", max_iter=120)
blob["joint_tuner|q0"] = q0
blob["joint_tuner|final"] = np.array([st_, gm_, sp_])
blob["joint_tuner|history"] = np.array([h["step_size"], h["gamma"], h["accept_rate"]], dtype=np.float64)

path = os.path.join(HERE, "reference_host_golden.npz")
np.savez_compressed(path, **blob)
print("wrote", path, os.path.getsize(path), "bytes;", len(names), "target cases")
