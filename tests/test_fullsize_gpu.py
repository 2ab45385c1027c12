"""BASELINE.json's full sizes, where the CPU oracle would take minutes: size-independent properties of the path.

  * idempotence: the same call twice gives the same bits (work queues, item claiming and atomics included);
  * shard invariance: two half-size calls with `chain_offset` reproduce the full call bit for bit — the multi-GPU
    contract, checked here on one GPU at the bench's 65,536 chains;
  * gamma = 0: the GRAHMC kernel is the HMC kernel bit for bit in the strict build (GRAHMC.py:194 with gamma 0 -> s = 1);
  * Metropolis bookkeeping: a rejected iteration repeats the previous row exactly, an accepted one does not, and the
    in-kernel accept counters agree with the stored rows (accept_count, accept_rate = count / S, fused sum);
  * NUTS: the fused gradient counter equals sum(2^depth - 1) over the stored tree depths (run_benchmarks.py:724).
"""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

D = 20


def _bench_inputs(C):
    from mcmc_b200.benchmarks import targets as T
    from mcmc_b200.random import normal
    tgt = T.rosenbrock(D, scale=0.1)
    q0 = torch.from_numpy(1.0 + 0.5 * normal(100, (C, D))).cuda()
    w = torch.linspace(0.5, 2.0, D, dtype=torch.float64, device="cuda")
    return tgt, q0, w


def _run(sampler, tag, q0, **kw):
    from mcmc_b200.runtime import run_sampler
    out = run_sampler(sampler, tag, q0, **kw)
    torch.cuda.synchronize()
    return out


def _same(a, b, keys, what):
    for k in keys:
        assert torch.equal(a[k], b[k]), (what, k)


@pytest.mark.parametrize("L", [16, 128])
def test_hmc_bench_shape_properties(gpu_lib, L):
    from mcmc_b200 import _abi
    C, S = 65536, 64
    tgt, q0, w = _bench_inputs(C)
    kw = dict(key=7, step_size=0.02, num_steps=L, num_samples=S, burn_in=0, inv_mass_matrix=w)
    keys = ("samples", "log_probs", "accept_count", "accept_rate", "final_q", "final_lp", "final_g")
    full = _run(_abi.SAMPLER_HMC, tgt.log_prob_fn, q0, **kw)
    assert torch.isfinite(full["samples"]).all()
    # idempotence
    again = _run(_abi.SAMPLER_HMC, tgt.log_prob_fn, q0, **kw)
    _same(full, again, keys, "idempotence")
    # shard invariance (different grid, different chunking of the work queue, same Philox counters)
    h = C // 2
    lo = _run(_abi.SAMPLER_HMC, tgt.log_prob_fn, q0[:h], chain_offset=0, **kw)
    hi = _run(_abi.SAMPLER_HMC, tgt.log_prob_fn, q0[h:], chain_offset=h, **kw)
    assert torch.equal(full["samples"][:, :h], lo["samples"]) and torch.equal(full["samples"][:, h:], hi["samples"])
    assert torch.equal(full["final_g"][h:], hi["final_g"]) and torch.equal(full["accept_count"][:h], lo["accept_count"])
    # gamma = 0 GRAHMC == HMC: in the reference's operation order (strict build; the fast build merges the friction
    # scaling and the kick into fma(p, A, g*B), one more rounding than HMC's fma(eps, g, p))
    kws = dict(kw, num_samples=8, variant="strict")
    h0 = _run(_abi.SAMPLER_HMC, tgt.log_prob_fn, q0, **kws)
    g0 = _run(_abi.SAMPLER_GRAHMC, tgt.log_prob_fn, q0, gamma=0.0, steepness=2.0, schedule_id=1, **kws)
    _same(h0, g0, keys, "gamma=0")
    # Metropolis bookkeeping
    smp = full["samples"]
    prev = torch.cat([q0.unsqueeze(0), smp[:-1]], 0)
    moved = (smp != prev).any(dim=2)                                  # (S, C): the row changed <=> accepted
    assert torch.equal(moved.sum(0).to(torch.int32), full["accept_count"])
    assert int(full["counters"][0]) == int(full["accept_count"].sum())
    assert int(full["counters"][1]) == C * S * L
    assert torch.equal(full["accept_rate"], full["accept_count"].float() / S)
    assert torch.equal(full["final_q"], smp[-1])
    acc = float(full["accept_rate"].mean())
    assert 0.5 < acc < 0.999, acc                                     # eps 0.02 on rosenbrock(0.1): ~0.8 (bench line)
    # log_probs are the target's values at the stored rows
    from mcmc_b200.runtime import target_eval
    lp, _ = target_eval(tgt.log_prob_fn, smp[-1])
    assert torch.allclose(lp, full["log_probs"][-1], rtol=1e-12, atol=1e-9)


def test_rwmh_four_million_chains_properties(gpu_lib):
    from mcmc_b200 import _abi
    from mcmc_b200.benchmarks import targets as T
    C, S = 4194304, 6                                                 # 4 GB of samples: config 4's chain count
    tgt = T.standard_normal(D)
    g = torch.Generator(device="cuda").manual_seed(5)
    q0 = 0.1 * torch.randn((C, D), dtype=torch.float64, device="cuda", generator=g)
    kw = dict(key=3, step_size=2.38 / np.sqrt(D), num_samples=S, burn_in=0)
    full = _run(_abi.SAMPLER_RWMH, tgt.log_prob_fn, q0, **kw)
    smp = full["samples"]
    prev = torch.cat([q0.unsqueeze(0), smp[:-1]], 0)
    moved = (smp != prev).any(dim=2)
    assert torch.equal(moved.sum(0).to(torch.int32), full["accept_count"])
    assert int(full["counters"][0]) == int(full["accept_count"].sum())
    assert torch.equal(full["final_q"], smp[-1])
    del prev, moved
    # the last quarter as its own shard
    q = 3 * C // 4
    part = _run(_abi.SAMPLER_RWMH, tgt.log_prob_fn, q0[q:], chain_offset=q, **kw)
    assert torch.equal(smp[:, q:], part["samples"]) and torch.equal(full["log_probs"][:, q:], part["log_probs"])
    lp_ref = -0.5 * (smp[-1] ** 2).sum(1) - 0.5 * D * np.log(2 * np.pi)
    assert torch.allclose(lp_ref, full["log_probs"][-1], rtol=1e-12, atol=1e-9)


@pytest.mark.parametrize("name,eps", [("rosenbrock", 0.02), ("neals_funnel", 0.1)])
def test_nuts_config5_shape_properties(gpu_lib, name, eps):
    from mcmc_b200 import _abi
    from mcmc_b200.benchmarks import targets as T
    from mcmc_b200.random import normal
    C, S = 65536, 24
    tgt = T.get_target(name, D)
    q0 = torch.from_numpy(1.0 + 0.5 * normal(300, (C, D)) if name == "rosenbrock" else tgt.init_sampler(300, C)).cuda()
    kw = dict(key=2, step_size=eps, num_samples=S, burn_in=0)
    keys = ("samples", "log_probs", "tree_depths", "mean_accept_probs", "final_q", "final_lp", "final_g", "accept_count")
    full = _run(_abi.SAMPLER_NUTS, tgt.log_prob_fn, q0, **kw)
    again = _run(_abi.SAMPLER_NUTS, tgt.log_prob_fn, q0, **kw)         # lanes claim different items: same bits
    _same(full, again, keys, "idempotence")
    dep = full["tree_depths"].to(torch.int64)
    assert int(full["counters"][1]) == int(((2 ** dep) - 1).sum())    # gradient evaluations
    assert int(full["counters"][2]) == int(dep.sum())
    assert int(dep.min()) >= 1 and int(dep.max()) <= 10
    assert torch.equal(full["accept_count"], torch.full((C,), S, dtype=torch.int32, device="cuda"))  # NUTS.py:512
    assert torch.equal(full["final_q"], full["samples"][-1])
    h = C // 2
    hi = _run(_abi.SAMPLER_NUTS, tgt.log_prob_fn, q0[h:], chain_offset=h, **kw)
    assert torch.equal(full["samples"][:, h:], hi["samples"]) and torch.equal(dep[:, h:], hi["tree_depths"].to(torch.int64))
    m = full["mean_accept_probs"]
    assert torch.isfinite(m).all() and float(m.min()) >= 0.0 and float(m.max()) <= 1.0
