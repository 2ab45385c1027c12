"""Oracle-vs-CUDA parity AT the BASELINE chain counts.

The runs below are the production path: native in-kernel Philox, the device-pointer C ABI, 65,536 / 4,194,304 / 32,768
chains, so the schedulers are sized as in production (HMC: iteration chunks of 450/(L+14) leapfrog-step equivalents
handed between persistent blocks; L >= 96: the 6-blocks-per-SM instantiation; NUTS: ~6 (chunk, chain) items per chain
claimed per lane; RWMH: warp tiles + per-warp TMA stores).  The oracle then replays 64-chain SLICES of each run — the
first chains, a window straddling block boundaries in the middle, and the last chains (ragged where C is not a multiple
of the block) — and every stored value of the slice is compared.

How a slice is replayed: the Philox stream is keyed on the GLOBAL chain id, so mcmc_philox_fill with chain_offset =
slice start returns exactly the normals the sampler kernel drew for these chains (tests/test_rng_gpu.py pins that stream
against an 80-bit Box-Muller); the accept / direction / proposal uniforms are k * 2^-53 with k an integer of the Philox
output — computed exactly in NumPy.  The oracle runs in injected mode on those draws.  With identical draws on both
sides the strict build (reference operation order, no FMA contraction) must reproduce the oracle BIT FOR BIT on targets
whose path has no libm call — also where the trajectories are chaotic (rosenbrock at the bench's eps = 0.02, L = 128: a
1-ulp perturbation grows to O(1) within eight iterations, so no tolerance could compare the fast build there).  The
fast build is compared at 1e-10 (HMC/NUTS: measured against the oracle's own response to a 1-ulp perturbation where
the dynamics amplify) on inputs whose trajectories are well-conditioned.
"""
import numpy as np
import pytest
import torch

from helpers import relerr, tid
from mcmc_b200 import _abi
from mcmc_b200.random import as_key, philox4x32_10

pytestmark = pytest.mark.gpu
D = 20


def _slices(C, n=64):
    mid = (C // 2 - n // 2) | 1                      # odd start: straddles 32- and 64-chain block boundaries
    return [(0, n), (mid, n), (C - n - 5, n + 5)]    # the last window ends at the last chain


def _u53(h, l):
    return ((h.astype(np.uint64) << np.uint64(21)) | (l.astype(np.uint64) >> np.uint64(11))).astype(np.float64) * 2.0 ** -53


def _native_draws(key, off, n, T, md, variant):
    """(z (T,n,D) as drawn on the device, u (T,n), dir_u (T,n,md), prop_u (T,n,md)) of global chains [off, off+n)."""
    from mcmc_b200.runtime import philox_fill
    k = as_key(key)
    z, _ = philox_fill(k, n, D, 0, T, chain_offset=off, variant=variant)
    gid = np.arange(n, dtype=np.uint64) + np.uint64(off)
    lo, hi = (gid & np.uint64(0xFFFFFFFF)).astype(np.uint32), (gid >> np.uint64(32)).astype(np.uint32)
    u = np.zeros((T, n))
    du = np.zeros((T, n, max(md, 1)))
    pu = np.zeros((T, n, max(md, 1)))
    for t in range(T):
        itv = np.full(n, t, dtype=np.uint32)
        o = philox4x32_10(lo, hi, itv, np.full(n, 1 << 24, dtype=np.uint32), k[0], k[1])
        u[t] = _u53(o[0], o[1])
        for d in range(md):
            o = philox4x32_10(lo, hi, itv, np.full(n, (2 << 24) | d, dtype=np.uint32), k[0], k[1])
            du[t, :, d] = _u53(o[0], o[1])
            pu[t, :, d] = _u53(o[2], o[3])
    return z.cpu().numpy(), u, du, pu


def _device_run(sampler, target, q0, variant, **kw):
    from mcmc_b200.runtime import run_sampler
    out = run_sampler(sampler, target.log_prob_fn, q0, variant=variant, **kw)
    torch.cuda.synchronize()
    return out


def _oracle_slice(oracle, sampler, tname, tpar, q0, off, n, key, S, B, L, eps, variant, w=None, md=0, flags=0, **extra):
    z, u, du, pu = _native_draws(key, off, n, B + S, md, variant)
    p = oracle.make_params(tid(tname), tpar, n, D, S, B, L, eps, rng_mode=_abi.RNG_INJECTED, flags=flags,
                           max_tree_depth=max(md, 1), **extra)
    kw = dict(z=z, u=u)
    if sampler == _abi.SAMPLER_NUTS:
        kw.update(dir_u=du, prop_u=pu)
    if flags & _abi.FLAG_ACCUM_WELFORD:
        kw["welford"] = (np.zeros((n, D)), np.zeros((n, D)))
    if sampler != _abi.SAMPLER_RWMH and w is not None:
        kw["inv_mass"] = w
    return oracle.run(sampler, p, q0[off:off + n].cpu().numpy(), store_samples=not (flags & _abi.FLAG_NO_SAMPLES), **kw)


def _close(a, b, exact, tol, tag):
    if exact:
        assert np.array_equal(a, b, equal_nan=True), tag
        return
    # a proposal that overflowed (rejected through the H1 = 1e10 guard) is non-finite on both sides, but inf vs nan
    # depends on the operation order: compare where the oracle is finite, require non-finite elsewhere
    fin = np.isfinite(b)
    assert np.array_equal(np.isfinite(a), fin), tag
    e = relerr(np.where(fin, a, 0.0), np.where(fin, b, 0.0))
    assert e <= tol, tag + (e, tol)


def _check(out, o, off, n, exact, tol, keys_sc=("samples", "log_probs"), keys_c=("final_q", "final_lp", "final_g"),
           what=""):
    for k in keys_sc:
        if k in out and k in o:
            _close(out[k][:, off:off + n].cpu().numpy(), o[k], exact, tol, (what, k, off))
    for k in keys_c:
        if k in out and k in o:
            _close(out[k][off:off + n].cpu().numpy(), o[k], exact, tol, (what, k, off))
    assert np.array_equal(out["accept_count"][off:off + n].cpu().numpy(), o["accept_count"]), (what, off)


def _per_chain_err(a, b, axis):
    """max over everything but the chain axis of |a - b| / max(|b|, 1), where both are finite (inf/nan patterns must agree)."""
    fin = np.isfinite(b)
    with np.errstate(invalid="ignore"):
        e = np.where(fin & np.isfinite(a), np.abs(a - b) / np.maximum(np.abs(b), 1.0), 0.0)
    e = np.moveaxis(e, axis, 0).reshape(e.shape[axis], -1).max(axis=1)
    bad = np.moveaxis(np.isfinite(a) != fin, axis, 0).reshape(e.shape[0], -1).any(axis=1)
    return e, bad


def _check_fast(got, o, o1, floor, what, max_chaotic=0.10):
    """Fast build vs oracle, chain by chain.  `got`: dict of arrays already cut to the slice; o / o1: the oracle on the
    slice and on the slice with the initial position moved by one ulp.  A chain's tolerance is max(floor, 100 x its own
    response to that ulp); chains whose response exceeds 1e-8 are chaotic over this run (no tolerance separates a
    reordered sum from a wrong result there) and are left out — at most `max_chaotic` of the slice — while every other
    chain must also reproduce its accept count exactly.  Returns the number of chains compared."""
    n = o["accept_count"].shape[0]
    sens = np.zeros(n)
    err = np.zeros(n)
    mism = np.zeros(n, dtype=bool)
    for k, arr in got.items():
        if k not in o or k == "accept_count":
            continue
        axis = 0 if k in ("final_q", "final_g", "final_lp", "welford_mean", "welford_m2") else 1   # the chain axis
        s_k, _ = _per_chain_err(o1[k], o[k], axis)
        e_k, b_k = _per_chain_err(arr, o[k], axis)
        sens = np.maximum(sens, s_k)
        err = np.maximum(err, e_k)
        mism |= b_k
    sens = np.maximum(sens, (o1["accept_count"] != o["accept_count"]) * 1.0)    # an ulp flipped a decision: chaotic
    ok = sens <= 1e-8
    assert ok.mean() >= 1.0 - max_chaotic, (what, "chaotic chains", int((~ok).sum()), n)
    tol = np.maximum(floor, 100.0 * sens)
    worst = np.argmax(np.where(ok, err / tol, 0.0))
    assert np.all(err[ok] <= tol[ok]), (what, "chain", int(worst), err[worst], tol[worst])
    assert not mism[ok].any(), (what, "finite pattern")
    assert np.array_equal(got["accept_count"][ok], o["accept_count"][ok]), what
    assert np.median(err) <= 100 * floor, (what, "median", float(np.median(err)))
    return int(ok.sum())


# ---------------------------------------------------------------------------------------------- config 2

@pytest.mark.parametrize("L,C", [(16, 65536), (16, 65500), (32, 65536), (64, 65536), (128, 65536), (128, 65500)])
def test_config2_hmc_rosenbrock_bench_inputs_strict_is_bit_exact(gpu_lib, oracle, L, C):
    """bench.py's inputs (init 1 + 0.5 N, eps 0.02, M^-1 = linspace(0.5, 2)), production chunking (3 chunks at L=16, 11 at
    L=128 where the 6-blocks-per-SM instantiation runs): the strict build equals the oracle bit for bit on every slice."""
    from mcmc_b200.benchmarks import targets as T
    from mcmc_b200.random import normal
    tgt = T.rosenbrock(D, scale=0.1)
    q0 = torch.from_numpy(1.0 + 0.5 * normal(100, (C, D))).cuda()
    w = np.linspace(0.5, 2.0, D)
    S, B, key = 24, 8, (11, L)
    out = _device_run(_abi.SAMPLER_HMC, tgt, q0, "strict", key=np.array(key, dtype=np.uint32), step_size=0.02,
                      num_steps=L, num_samples=S, burn_in=B, inv_mass_matrix=w)
    assert int(out["counters"][1]) == C * (S + B) * L
    for off, n in _slices(C):
        o = _oracle_slice(oracle, _abi.SAMPLER_HMC, "rosenbrock", [0.1], q0, off, n, key, S, B, L, 0.02, "strict", w=w)
        _check(out, o, off, n, exact=True, tol=0.0, what=f"strict L={L}")


@pytest.mark.parametrize("L,eps,spread,tol", [(16, 0.02, 0.5, 1e-9), (32, 0.005, 0.05, 1e-10), (64, 0.005, 0.05, 1e-10),
                                               (128, 0.005, 0.05, 1e-10)])
def test_config2_hmc_rosenbrock_fast_build_slices(gpu_lib, oracle, L, eps, spread, tol):
    """The product (fast) build at 65,536 chains.  L = 16 on the bench's inputs (a 1-ulp perturbation of the initial
    position moves a 32-iteration run by 8e-11 there; at L = 32 already by 3e-6, at L = 128 by O(1): those are covered bit
    for bit by the strict build above); L = 32 / 64 / 128 — production chunking and the 6-blocks-per-SM instantiation — on
    well-conditioned inputs (init 1 + 0.05 N, eps 0.005).  Yardstick: 1e-10, or 100x the oracle's own response to a 1-ulp
    perturbation of the initial position (over all compared outputs) where that is larger."""
    from mcmc_b200.benchmarks import targets as T
    from mcmc_b200.random import normal
    C, S, B, key = 65536, 24, 8, (12, L)
    tgt = T.rosenbrock(D, scale=0.1)
    q0 = torch.from_numpy(1.0 + spread * normal(101, (C, D))).cuda()
    w = np.linspace(0.5, 2.0, D)
    out = _device_run(_abi.SAMPLER_HMC, tgt, q0, "fast", key=np.array(key, dtype=np.uint32), step_size=eps,
                      num_steps=L, num_samples=S, burn_in=B, inv_mass_matrix=w, track_proposals=True)
    for off, n in _slices(C):
        o = _oracle_slice(oracle, _abi.SAMPLER_HMC, "rosenbrock", [0.1], q0, off, n, key, S, B, L, eps, "fast", w=w,
                          flags=_abi.FLAG_TRACK_PROPOSALS)
        q1 = q0.clone()
        q1[off:off + n] = torch.from_numpy(np.nextafter(q0[off:off + n].cpu().numpy(), np.inf)).cuda()
        o1 = _oracle_slice(oracle, _abi.SAMPLER_HMC, "rosenbrock", [0.1], q1, off, n, key, S, B, L, eps, "fast", w=w,
                           flags=_abi.FLAG_TRACK_PROPOSALS)
        keys = ("samples", "log_probs", "prop_q", "prop_lp", "delta_h", "pre_q")
        got = {k: out[k][:, off:off + n].cpu().numpy() for k in keys}
        got.update({k: out[k][off:off + n].cpu().numpy() for k in ("final_q", "final_lp", "final_g", "accept_count")})
        _check_fast(got, o, o1, tol, f"fast L={L} off={off}")


# ---------------------------------------------------------------------------------------------- config 3 (one tuning batch)

@pytest.mark.parametrize("sched,steep", [("tanh", 0.5), ("linear", 2.0)])
def test_config3_grahmc_funnel_tuning_batch_slices(gpu_lib, oracle, sched, steep):
    """One 100-iteration adaptation batch of run_adaptive_warmup at 32,768 chains per GPU: GRAHMC on Neal's funnel from
    the target's own init sampler, samples not stored, per-chain Welford accumulated in the kernel's shared-memory tiles."""
    from mcmc_b200.benchmarks import targets as T
    C, S, L, eps, key = 32768, 100, 16, 0.05, (13, 1)
    tgt = T.neals_funnel(D)
    q0 = torch.as_tensor(np.asarray(tgt.init_sampler(5, C)), dtype=torch.float64).cuda()
    wm = torch.zeros((C, D), dtype=torch.float64, device="cuda")
    w2 = torch.zeros_like(wm)
    out = _device_run(_abi.SAMPLER_GRAHMC, tgt, q0, "fast", key=np.array(key, dtype=np.uint32), step_size=eps,
                      num_steps=L, num_samples=S, gamma=1.0, steepness=steep, schedule_id=_abi.SCHEDULE_IDS[sched],
                      welford=(wm, w2), store_samples=False)
    fl = _abi.FLAG_ACCUM_WELFORD | _abi.FLAG_NO_SAMPLES
    for off, n in _slices(C):
        o = _oracle_slice(oracle, _abi.SAMPLER_GRAHMC, "neals_funnel", [], q0, off, n, key, S, 0, L, eps, "fast",
                          flags=fl, gamma=1.0, steepness=steep, schedule_id=_abi.SCHEDULE_IDS[sched])
        q1 = q0.clone()
        q1[off:off + n] = torch.from_numpy(np.nextafter(q0[off:off + n].cpu().numpy(), np.inf)).cuda()
        o1 = _oracle_slice(oracle, _abi.SAMPLER_GRAHMC, "neals_funnel", [], q1, off, n, key, S, 0, L, eps, "fast",
                           flags=fl, gamma=1.0, steepness=steep, schedule_id=_abi.SCHEDULE_IDS[sched])
        got = {k: out[k][off:off + n].cpu().numpy() for k in ("final_q", "final_lp", "final_g", "accept_count")}
        got["welford_mean"] = wm[off:off + n].cpu().numpy()
        got["welford_m2"] = w2[off:off + n].cpu().numpy()
        # 100 iterations on the funnel: chains near the neck amplify an ulp beyond any tolerance (left out, <= 25%)
        assert _check_fast(got, o, o1, 1e-10, f"{sched} off={off}", max_chaotic=0.25) >= 48
    assert int(out["counters"][0]) == int(out["accept_count"].sum())


# ---------------------------------------------------------------------------------------------- config 4

@pytest.mark.parametrize("variant,C", [("fast", 4194304), ("fast", 4194304 - 37), ("strict", 4194304)])
def test_config4_rwmh_4m_chains_slices(gpu_lib, oracle, variant, C):
    """RWMH on the 20-D standard normal at 4 M chains (scale 2.38/sqrt(20), init 0.1 N): warp tiles, per-warp bulk stores;
    the strict build bit for bit, the fast build at 1e-10; C - 37 leaves a ragged last warp (plain-store tail)."""
    from mcmc_b200.benchmarks import targets as T
    S, B, key, scale = 6, 2, (14, 2), 2.38 / np.sqrt(D)
    tgt = T.standard_normal(D)
    g = torch.Generator(device="cuda").manual_seed(3)
    q0 = 0.1 * torch.randn((C, D), dtype=torch.float64, device="cuda", generator=g)
    out = _device_run(_abi.SAMPLER_RWMH, tgt, q0, variant, key=np.array(key, dtype=np.uint32), step_size=scale,
                      num_samples=S, burn_in=B)
    for off, n in _slices(C):
        o = _oracle_slice(oracle, _abi.SAMPLER_RWMH, "standard_normal", [1.0], q0, off, n, key, S, B, 0, scale, variant)
        _check(out, o, off, n, exact=variant == "strict", tol=1e-10, what=f"rwmh {variant}")
    acc = float(out["accept_rate"].mean())
    assert 0.03 < acc < 0.6, acc                    # 0.10 over the first 8 iterations from the 0.1 N start (0.25 once spread out)


# ---------------------------------------------------------------------------------------------- config 5

@pytest.mark.parametrize("tname,tpar,eps,variant,C", [
    ("rosenbrock", [0.1], 0.03, "strict", 65536), ("rosenbrock", [0.1], 0.03, "strict", 65500),
    ("neals_funnel", [], 0.15, "fast", 65536), ("standard_normal", [1.0], 0.3, "fast", 65536)])
def test_config5_nuts_65536_chains_slices(gpu_lib, oracle, tname, tpar, eps, variant, C):
    """NUTS, max_tree_depth 10, at 65,536 chains: more chains than resident lanes, so the (chunk, chain) item queue and
    the home-row hand-off run as in production.  Tree depths identical; rosenbrock (no libm on the trajectory) in the
    strict build bit for bit from the target's own init distribution; funnel / standard normal in the fast build at 1e-10."""
    from mcmc_b200.benchmarks import targets as T
    S, B, md, key = 8, 4, 10, (15, len(tname))
    tgt = {"rosenbrock": lambda: T.rosenbrock(D, scale=0.1), "neals_funnel": lambda: T.neals_funnel(D),
           "standard_normal": lambda: T.standard_normal(D)}[tname]()
    if tgt.init_sampler is not None:
        q0 = torch.as_tensor(np.asarray(tgt.init_sampler(6, C)), dtype=torch.float64).cuda()
    else:
        q0 = torch.randn((C, D), dtype=torch.float64, device="cuda", generator=torch.Generator(device="cuda").manual_seed(4))
    out = _device_run(_abi.SAMPLER_NUTS, tgt, q0, variant, key=np.array(key, dtype=np.uint32), step_size=eps,
                      num_samples=S, burn_in=B, max_tree_depth=md)
    exact = variant == "strict"
    for off, n in _slices(C):
        o = _oracle_slice(oracle, _abi.SAMPLER_NUTS, tname, tpar, q0, off, n, key, S, B, 0, eps, variant, md=md)
        assert np.array_equal(out["tree_depths"][:, off:off + n].cpu().numpy(), o["tree_depths"]), (tname, off)
        if exact:
            _check(out, o, off, n, exact=True, tol=0.0, what=f"nuts {tname}")
            # the Metropolis weights go through exp(): CUDA's and glibc's differ by an ulp, so not bit for bit
            assert relerr(out["mean_accept_probs"][:, off:off + n].cpu().numpy(), o["mean_accept_probs"]) <= 1e-12
            continue
        q1 = q0.clone()
        q1[off:off + n] = torch.from_numpy(np.nextafter(q0[off:off + n].cpu().numpy(), np.inf)).cuda()
        o1 = _oracle_slice(oracle, _abi.SAMPLER_NUTS, tname, tpar, q1, off, n, key, S, B, 0, eps, variant, md=md)
        got = {k: out[k][:, off:off + n].cpu().numpy() for k in ("samples", "log_probs", "mean_accept_probs")}
        got.update({k: out[k][off:off + n].cpu().numpy() for k in ("final_q", "final_lp", "final_g", "accept_count")})
        # depth-10 trees (1023 steps) in the funnel's neck amplify an ulp: those chains are left out (<= 25%)
        _check_fast(got, o, o1, 1e-10, f"nuts {tname} off={off}", max_chaotic=0.25)
    d = out["tree_depths"].to(torch.int64)
    assert int(out["counters"][1]) >= int(((1 << d) - 1).sum())       # burn-in transitions count too
    assert int(d.max()) <= md and int(d.min()) >= 1
