"""Generates tests/golden/reference_quickstart_golden.npz: BASELINE.json configs[0] — the reference's README quickstart
(README.md:41-71: GRAHMC tanh, 10-D `-0.5*sum(x**2)`, 4 chains, 1000 samples after 500 burn-in, step 0.05, L=16,
gamma 1, steepness 0.5, identity metric) — executed AT FULL SIZE by the reference's own samplers/GRAHMC.py on the
torch-fp64 stand-in for jax (see make_reference_golden.py).  Stored: the initial positions, all 1500 iterations of
recorded momenta / uniforms, every log-prob, every 20th stored position, the final state and the accept counts.
Build container only (about a minute)."""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
import make_reference_golden as G  # noqa: E402
import jax.numpy as jnp  # noqa: E402
from jax import random  # noqa: E402
from jax._core import DRAW_LOG  # noqa: E402


def log_prob_fn(x):                      # README.md:51-52, verbatim
    return -0.5 * jnp.sum(x**2, axis=-1)


key = random.PRNGKey(0)
init_position = random.normal(key, shape=(4, 10)) * 2.0
del DRAW_LOG[:]
samples, log_probs, accept_rate, final_state = G.GRAHMC.rahmc_run(
    key, log_prob_fn, init_position,
    num_samples=1000,
    burn_in=500,
    step_size=0.05,
    num_steps=16,
    gamma=1.0,
    steepness=0.5,
    friction_schedule=G.GRAHMC.get_friction_schedule('tanh'))
log = list(DRAW_LOG)
assert len(log) == 3000
np64 = G.np64
blob = dict(
    q0=np64(init_position),
    z=np.stack([log[2 * i][2].reshape(4, 10) for i in range(1500)]),
    u=np.stack([log[2 * i + 1][2].reshape(4) for i in range(1500)]),
    log_probs=np64(log_probs), samples_every20=np64(samples)[::20].copy(), samples_last=np64(samples)[-1],
    accept_rate=np64(accept_rate).astype(np.float32), accept_count=np64(final_state.accept_count).astype(np.int32),
    final_q=np64(final_state.position), final_lp=np64(final_state.log_prob), final_g=np64(final_state.grad_log_prob),
    sample_mean=np64(samples).mean(axis=(0, 1)), sample_var=np64(samples).var(axis=(0, 1)))
path = os.path.join(HERE, "reference_quickstart_golden.npz")
np.savez_compressed(path, **blob)
print("accept", float(np64(accept_rate).mean()), "wrote", path, os.path.getsize(path), "bytes")
