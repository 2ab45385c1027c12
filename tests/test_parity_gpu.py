"""GPU parity: the CUDA path (through the C ABI) against the CPU oracle on the same seeded
inputs with injected momenta / uniforms (north_star, correctness check 1).

Tolerances (written here, stated in DESIGN.md):
  * strict build (-DMCMC_FAITHFUL -fmad=false): BIT-EXACT on targets whose path has no libm
    call (standard_normal, ill_conditioned, rosenbrock), <= 1e-12 elsewhere (exp/log/tanh of
    CUDA and glibc differ by <= 1-2 ulp);
  * fast build (FMA contraction, merged half-kicks): every leapfrog endpoint, log-prob and
    Hamiltonian within 1e-10 relative (|a-b| / max(|b|, 1)), every accept decision identical.
"""
import numpy as np
import pytest

from golden.make_golden import CASES, run_case
from helpers import POLYNOMIAL, TARGET_CASES, GpuBackend, relerr, tid
from mcmc_b200 import _abi

pytestmark = pytest.mark.gpu

SCHEDS = list(_abi.SCHEDULE_IDS)
TRACK = ("samples", "log_probs", "pre_q", "pre_lp", "prop_q", "prop_lp", "delta_h", "final_q", "final_lp", "final_g")


def _inputs(name, C_, D, S, B, seed, md=8):
    params, eps, init = TARGET_CASES[name]
    rng = np.random.default_rng(seed)
    return dict(params=params, eps=eps, q0=init(rng, C_, D), z=rng.standard_normal((B + S, C_, D)),
                u=rng.random((B + S, C_)), du=rng.random((B + S, C_, md)), pu=rng.random((B + S, C_, md)),
                w=np.linspace(0.5, 2.0, D))


def _compare(r, o, keys, tol, exact=False):
    for k in keys:
        if exact:
            assert np.array_equal(r[k], o[k], equal_nan=True), k
        else:
            assert relerr(r[k], o[k]) <= tol, (k, relerr(r[k], o[k]))
    assert np.array_equal(r["accept_count"], o["accept_count"])
    assert np.array_equal(r["accept_rate"], o["accept_rate"])
    assert int(r["counters"][0]) == int(o["counters"][0]) and int(r["counters"][1]) == int(o["counters"][1])


@pytest.mark.parametrize("variant", ["strict", "fast"])
@pytest.mark.parametrize("D", [20, 10, 7])
@pytest.mark.parametrize("name", list(TARGET_CASES))
def test_hmc_parity(gpu_lib, oracle, name, D, variant):
    C_, S, B, L = 100, 5, 2, 16   # 100 chains: one full 64-thread block plus a ragged tail
    i = _inputs(name, C_, D, S, B, seed=21)
    p = oracle.make_params(tid(name), i["params"], C_, D, S, B, L, i["eps"], flags=_abi.FLAG_TRACK_PROPOSALS)
    o = oracle.run(_abi.SAMPLER_HMC, p, i["q0"], inv_mass=i["w"], z=i["z"], u=i["u"])
    r = GpuBackend(variant).run(_abi.SAMPLER_HMC, p, i["q0"], inv_mass=i["w"], z=i["z"], u=i["u"])
    exact = variant == "strict" and name in POLYNOMIAL
    _compare(r, o, TRACK, 1e-12 if variant == "strict" else 1e-10, exact=exact)


@pytest.mark.parametrize("variant", ["strict", "fast"])
@pytest.mark.parametrize("sched", SCHEDS)
@pytest.mark.parametrize("name", ["neals_funnel", "standard_normal", "rosenbrock", "student_t"])
def test_grahmc_parity(gpu_lib, oracle, name, sched, variant):
    C_, D, S, B, L = 70, 20, 5, 2, 16
    i = _inputs(name, C_, D, S, B, seed=22)
    steep = 0.5 if sched == "tanh" else 2.0
    p = oracle.make_params(tid(name), i["params"], C_, D, S, B, L, i["eps"], gamma=1.0, steepness=steep,
                           schedule_id=_abi.SCHEDULE_IDS[sched], flags=_abi.FLAG_TRACK_PROPOSALS)
    o = oracle.run(_abi.SAMPLER_GRAHMC, p, i["q0"], inv_mass=i["w"], z=i["z"], u=i["u"])
    r = GpuBackend(variant).run(_abi.SAMPLER_GRAHMC, p, i["q0"], inv_mass=i["w"], z=i["z"], u=i["u"])
    _compare(r, o, TRACK, 1e-12 if variant == "strict" else 1e-10)


@pytest.mark.parametrize("L", [1, 2, 32, 64, 128])
def test_hmc_parity_L_grid_rosenbrock(gpu_lib, oracle, L):
    # BASELINE config 2's L grid on the chaotic target: tolerance 1e-10 at L <= 32; at L = 64/128 the
    # oracle's own sensitivity to a 1-ulp perturbation is reported and used as the yardstick.
    C_, D, S, B = 64, 20, 3, 1
    i = _inputs("rosenbrock", C_, D, S, B, seed=23)
    p = oracle.make_params(tid("rosenbrock"), i["params"], C_, D, S, B, L, i["eps"], flags=_abi.FLAG_TRACK_PROPOSALS)
    o = oracle.run(_abi.SAMPLER_HMC, p, i["q0"], z=i["z"], u=i["u"])
    r = GpuBackend("fast").run(_abi.SAMPLER_HMC, p, i["q0"], z=i["z"], u=i["u"])
    q1 = np.nextafter(i["q0"], np.inf)
    o1 = oracle.run(_abi.SAMPLER_HMC, p, q1, z=i["z"], u=i["u"])
    sens = max(relerr(o1[k], o[k]) for k in ("prop_q", "delta_h"))
    err = max(relerr(r[k], o[k]) for k in ("prop_q", "prop_lp", "delta_h", "samples"))
    tol = 1e-10 if L <= 32 else max(1e-10, 50 * sens)
    assert err <= tol, (L, err, sens)
    assert np.array_equal(r["accept_count"], o["accept_count"])
    rs = GpuBackend("strict").run(_abi.SAMPLER_HMC, p, i["q0"], z=i["z"], u=i["u"])
    for k in TRACK:
        assert np.array_equal(rs[k], o[k]), (L, k)  # strict build: bit-exact at every L


@pytest.mark.parametrize("variant", ["strict", "fast"])
@pytest.mark.parametrize("D", [20, 10, 3])
@pytest.mark.parametrize("name", list(TARGET_CASES))
def test_rwmh_parity(gpu_lib, oracle, name, D, variant):
    C_, S, B = 300, 12, 3
    i = _inputs(name, C_, D, S, B, seed=24)
    scale = 0.02 if name == "rosenbrock" else 2.38 / np.sqrt(D) * 0.5
    p = oracle.make_params(tid(name), i["params"], C_, D, S, B, 0, scale)
    o = oracle.run(_abi.SAMPLER_RWMH, p, i["q0"], z=i["z"], u=i["u"])
    r = GpuBackend(variant).run(_abi.SAMPLER_RWMH, p, i["q0"], z=i["z"], u=i["u"])
    exact = variant == "strict" and name in POLYNOMIAL
    for k in ("samples", "log_probs", "final_q", "final_lp"):
        if exact:
            assert np.array_equal(r[k], o[k]), k
        else:
            assert relerr(r[k], o[k]) <= 1e-10, k
    assert np.array_equal(r["accept_count"], o["accept_count"])


@pytest.mark.parametrize("variant", ["strict", "fast"])
@pytest.mark.parametrize("D", [20, 10, 6])
@pytest.mark.parametrize("name", ["standard_normal", "neals_funnel", "rosenbrock", "gaussian_mixture", "log_gamma"])
def test_nuts_parity(gpu_lib, oracle, name, D, variant):
    C_, S, B, md = 50, 4, 1, 8
    i = _inputs(name, C_, D, S, B, seed=25, md=md)
    p = oracle.make_params(tid(name), i["params"], C_, D, S, B, 0, i["eps"], max_tree_depth=md)
    kw = dict(inv_mass=i["w"], z=i["z"], u=i["u"], dir_u=i["du"], prop_u=i["pu"])
    o = oracle.run(_abi.SAMPLER_NUTS, p, i["q0"], **kw)
    r = GpuBackend(variant).run(_abi.SAMPLER_NUTS, p, i["q0"], **kw)
    assert np.array_equal(r["tree_depths"], o["tree_depths"])
    for k in ("samples", "log_probs", "mean_accept_probs", "final_q", "final_lp", "final_g"):
        assert relerr(r[k], o[k]) <= 1e-10, (k, relerr(r[k], o[k]))
    for j in range(4):
        assert int(r["counters"][j]) == int(o["counters"][j]), j


@pytest.mark.parametrize("case", CASES, ids=[c["id"] for c in CASES])
def test_golden_fixtures_gpu(gpu_lib, case):
    import os
    gold = np.load(os.path.join(os.path.dirname(__file__), "golden", "hot_path_golden.npz"))
    got = run_case(GpuBackend("fast"), case)
    for k, v in got.items():
        want = gold[f"{case['id']}/{k}"]
        if v.dtype.kind in "iu":
            assert np.array_equal(v, want), k
        else:
            assert relerr(v, want) <= 1e-10, (k, relerr(v, want))


def test_edge_cases(gpu_lib, oracle):
    g = GpuBackend("fast")
    # zero samples, zero chains, L = 0, C = 1 (the reference promotes a 1-D init to one chain, HMC.py:52-53)
    q1 = np.full((1, 20), 0.3)
    for S, B, L, Cn in [(0, 2, 4, 1), (3, 0, 0, 1), (2, 1, 3, 1)]:
        z = np.random.default_rng(1).standard_normal((B + S, Cn, 20))
        u = np.random.default_rng(2).random((B + S, Cn))
        p = oracle.make_params(tid("standard_normal"), [1.0], Cn, 20, S, B, L, 0.1)
        o = oracle.run(_abi.SAMPLER_HMC, p, q1, z=z, u=u)
        r = g.run(_abi.SAMPLER_HMC, p, q1, z=z, u=u)
        assert relerr(r["final_q"], o["final_q"]) <= 1e-12 and r["samples"].shape == (S, Cn, 20)
        assert relerr(r["samples"], o["samples"]) <= 1e-12
    p0 = oracle.make_params(tid("standard_normal"), [1.0], 0, 20, 2, 0, 4, 0.1, rng_mode=_abi.RNG_PHILOX)
    r = g.run(_abi.SAMPLER_HMC, p0, np.zeros((0, 20)))
    assert r["samples"].shape == (2, 0, 20)
    # non-finite proposal: H1 -> 1e10 -> rejected (HMC.py:199); log_gamma pushed negative -> lp = -inf
    C_, D = 40, 20
    rng = np.random.default_rng(9)
    q0 = 0.05 + 0.01 * rng.random((C_, D))
    z = 5.0 * rng.standard_normal((4, C_, D))
    u = rng.random((4, C_))
    p = oracle.make_params(tid("log_gamma"), TARGET_CASES["log_gamma"][0], C_, D, 4, 0, 8, 0.2,
                           flags=_abi.FLAG_TRACK_PROPOSALS)
    o = oracle.run(_abi.SAMPLER_HMC, p, q0, z=z, u=u)
    r = g.run(_abi.SAMPLER_HMC, p, q0, z=z, u=u)
    assert np.array_equal(r["accept_count"], o["accept_count"])
    assert np.array_equal(np.isinf(r["prop_lp"]), np.isinf(o["prop_lp"])) and np.isinf(o["prop_lp"]).any()
    assert relerr(r["samples"], o["samples"]) <= 1e-10


def test_welford_and_no_samples_flags(gpu_lib, oracle):
    C_, D, S = 90, 20, 7
    i = _inputs("neals_funnel", C_, D, S, 0, seed=31)
    fl = _abi.FLAG_ACCUM_WELFORD | _abi.FLAG_NO_SAMPLES
    p = oracle.make_params(tid("neals_funnel"), [], C_, D, S, 0, 8, i["eps"], flags=fl)
    wf = (np.zeros((C_, D)), np.zeros((C_, D)))
    o = oracle.run(_abi.SAMPLER_HMC, p, i["q0"], z=i["z"], u=i["u"], welford=wf)
    r = GpuBackend("fast").run(_abi.SAMPLER_HMC, p, i["q0"], z=i["z"], u=i["u"], welford=wf)
    assert relerr(r["welford_mean"], o["welford_mean"]) <= 1e-10
    assert relerr(r["welford_m2"], o["welford_m2"]) <= 1e-10
    assert "samples" not in r


def test_native_philox_stream_matches_oracle_and_is_shard_invariant(gpu_lib, oracle):
    # native RNG: same counter layout on both sides (oracle/rng_common.h), so whole runs agree to
    # libm-level differences; and chain_offset makes a shard reproduce its slice of the full run.
    C_, D, S, B, L = 96, 20, 6, 2, 8
    q0 = np.random.default_rng(5).standard_normal((C_, D))
    mk = lambda C2, off: oracle.make_params(tid("standard_normal"), [1.0], C2, D, S, B, L, 0.2,
                                            rng_mode=_abi.RNG_PHILOX, seed=(123, 456), chain_offset=off)
    g = GpuBackend("fast")
    full = g.run(_abi.SAMPLER_HMC, mk(C_, 0), q0)
    o = oracle.run(_abi.SAMPLER_HMC, mk(C_, 0), q0)
    assert relerr(full["samples"], o["samples"]) <= 1e-9
    assert np.array_equal(full["accept_count"], o["accept_count"])
    half = g.run(_abi.SAMPLER_HMC, mk(C_ // 2, C_ // 2), q0[C_ // 2:])
    assert np.array_equal(half["samples"], full["samples"][:, C_ // 2:])
    for smp, extra in ((_abi.SAMPLER_RWMH, {}), (_abi.SAMPLER_NUTS, {})):
        pf = mk(C_, 0)
        a = g.run(smp, pf, q0)
        b = oracle.run(smp, pf, q0)
        assert relerr(a["samples"], b["samples"]) <= 1e-8, smp


@pytest.mark.parametrize("sampler,sched", [(_abi.SAMPLER_HMC, None), (_abi.SAMPLER_GRAHMC, "tanh")])
def test_chunked_work_queue_is_bit_identical(gpu_lib, oracle, sampler, sched, monkeypatch):
    """The wave-balanced scheduler (iteration chunks handed between blocks through HBM) must not change
    a single bit: counters are keyed on (chain, absolute iteration), the state hand-off is exact."""
    C_, D, S, B, L = 300, 20, 11, 3, 8
    i = _inputs("neals_funnel", C_, D, S, B, seed=41)
    fl = _abi.FLAG_TRACK_PROPOSALS | _abi.FLAG_ACCUM_WELFORD
    mk = lambda mode: oracle.make_params(tid("neals_funnel"), [], C_, D, S, B, L, i["eps"], gamma=1.0, steepness=0.5,
                                         schedule_id=_abi.SCHEDULE_IDS.get(sched, 0), rng_mode=mode, seed=(7, 9),
                                         flags=fl)
    wf = (np.zeros((C_, D)), np.zeros((C_, D)))
    g = GpuBackend("fast")
    for mode, kw in ((_abi.RNG_INJECTED, dict(z=i["z"], u=i["u"])), (_abi.RNG_PHILOX, {})):
        monkeypatch.delenv("MCMC_HMC_FORCE_CHUNKS", raising=False)
        ref = g.run(sampler, mk(mode), i["q0"], inv_mass=i["w"], welford=wf, **kw)
        for nchunks in ("2", "5", "14"):
            monkeypatch.setenv("MCMC_HMC_FORCE_CHUNKS", nchunks)
            got = g.run(sampler, mk(mode), i["q0"], inv_mass=i["w"], welford=wf, **kw)
            for k in ref:
                assert np.array_equal(ref[k], got[k], equal_nan=True), (mode, nchunks, k)
    monkeypatch.delenv("MCMC_HMC_FORCE_CHUNKS", raising=False)
    o = oracle.run(sampler, mk(_abi.RNG_INJECTED), i["q0"], inv_mass=i["w"], z=i["z"], u=i["u"], welford=wf)
    monkeypatch.setenv("MCMC_HMC_FORCE_CHUNKS", "4")
    got = g.run(sampler, mk(_abi.RNG_INJECTED), i["q0"], inv_mass=i["w"], z=i["z"], u=i["u"], welford=wf)
    for k in ("samples", "prop_q", "delta_h", "final_g", "welford_mean", "welford_m2"):
        assert relerr(got[k], o[k]) <= 1e-10, k
    assert np.array_equal(got["accept_count"], o["accept_count"])


def test_nuts_native_stream_segments_are_bit_identical(gpu_lib, oracle, monkeypatch):
    """NUTS draws its momenta with a separate convergent kernel into a bounded buffer; a run longer than the buffer is
    cut into segments of iterations whose state travels through the workspace.  Neither must change a bit, and the
    stream equals the oracle's in-loop Philox draws."""
    C_, D, S, B = 70, 20, 9, 3
    q0 = np.random.default_rng(8).standard_normal((C_, D))
    wf = (np.zeros((C_, D)), np.zeros((C_, D)))
    p = oracle.make_params(tid("neals_funnel"), [], C_, D, S, B, 0, 0.15, rng_mode=_abi.RNG_PHILOX, seed=(5, 6),
                           max_tree_depth=6, flags=_abi.FLAG_ACCUM_WELFORD)
    g = GpuBackend("fast")
    monkeypatch.delenv("MCMC_NUTS_RNG_CAP_MB", raising=False)
    ref = g.run(_abi.SAMPLER_NUTS, p, q0, welford=wf)
    o = oracle.run(_abi.SAMPLER_NUTS, p, q0, welford=wf)
    assert np.array_equal(ref["tree_depths"], o["tree_depths"])
    for k in ("samples", "log_probs", "mean_accept_probs", "final_q", "final_g", "welford_mean", "welford_m2"):
        assert relerr(ref[k], o[k]) <= 1e-9, k
    for j in range(4):
        assert int(ref["counters"][j]) == int(o["counters"][j]), j
    # 70 chains x 21 doubles = 11.8 KB per iteration: a "0 MB" cap gives one-iteration segments (12 launches)
    monkeypatch.setenv("MCMC_NUTS_RNG_CAP_MB", "0")
    seg = g.run(_abi.SAMPLER_NUTS, p, q0, welford=wf)
    for k in ref:
        assert np.array_equal(ref[k], seg[k], equal_nan=True), k


@pytest.mark.parametrize("name,D,eps", [("rosenbrock", 20, 0.03), ("neals_funnel", 20, 0.15), ("gaussian_mixture", 7, 0.2)])
def test_nuts_items_claimed_per_lane_are_bit_identical(gpu_lib, oracle, monkeypatch, name, D, eps):
    """A NUTS chain is not tied to a lane: (chunk of iterations, chain) items are claimed from a queue and the state
    travels through the workspace, with lanes parked in a WAIT state while a predecessor item runs elsewhere.  Whatever
    the chunk length (1 iteration: every transition changes hands; 4: ragged last chunk; default: one item per chain at
    this size), compiled-D (shared-memory lane tiles) or runtime-D (L2 lane tiles), the outputs must not change a bit."""
    C_, S, B = 300, 11, 2
    q0 = np.random.default_rng(9).standard_normal((C_, D))
    wf = (np.zeros((C_, D)), np.zeros((C_, D)))
    par = [3.0] if name == "gaussian_mixture" else []
    p = oracle.make_params(tid(name), par, C_, D, S, B, 0, eps, rng_mode=_abi.RNG_PHILOX, seed=(15, 16),
                           max_tree_depth=7, flags=_abi.FLAG_ACCUM_WELFORD)
    g = GpuBackend("fast")
    monkeypatch.delenv("MCMC_NUTS_CHUNK", raising=False)
    ref = g.run(_abi.SAMPLER_NUTS, p, q0, welford=wf)
    o = oracle.run(_abi.SAMPLER_NUTS, p, q0, welford=wf)
    assert np.array_equal(ref["tree_depths"], o["tree_depths"])
    for j in range(4):
        assert int(ref["counters"][j]) == int(o["counters"][j]), j
    for ci in ("1", "4"):
        monkeypatch.setenv("MCMC_NUTS_CHUNK", ci)
        got = g.run(_abi.SAMPLER_NUTS, p, q0, welford=wf)
        for k in ref:
            assert np.array_equal(ref[k], got[k], equal_nan=True), (ci, k)


@pytest.mark.parametrize("name,sampler,sched,D", [("rosenbrock", _abi.SAMPLER_HMC, None, 20),
                                                  ("neals_funnel", _abi.SAMPLER_GRAHMC, "tanh", 20),
                                                  ("standard_normal", _abi.SAMPLER_GRAHMC, "constant", 10),
                                                  ("student_t", _abi.SAMPLER_HMC, None, 10)])
def test_warp_specialised_kernel_is_bit_identical(gpu_lib, oracle, monkeypatch, name, sampler, sched, D):
    """hmc_ws_kernel (producer warps draw the momenta into a shared-memory ring, consumer warps integrate) against
    hmc_kernel: same Philox counters, same arithmetic -> the same bits; with and without the chunked queue, with a
    ragged tail (C not a multiple of 32), Welford accumulation and the no-samples mode."""
    if b"+ws" not in gpu_lib.mcmc_build_info():
        pytest.skip("the warp-specialised kernel is an experiment compiled only with -DMCMC_WITH_WS "
                    "(measured 4-6% slower, profiles/ncu_hmc_ws_r01.txt)")
    C_, S, B, L = 700, 7, 2, 5
    i = _inputs(name, C_, D, S, B, seed=77)
    g = GpuBackend("fast")
    wf = (np.zeros((C_, D)), np.zeros((C_, D)))
    for flags in (_abi.FLAG_ACCUM_WELFORD, _abi.FLAG_NO_SAMPLES | _abi.FLAG_ACCUM_WELFORD, 0):
        p = oracle.make_params(tid(name), i["params"], C_, D, S, B, L, i["eps"], gamma=0.6, steepness=0.5,
                               schedule_id=_abi.SCHEDULE_IDS.get(sched, 0), rng_mode=_abi.RNG_PHILOX, seed=(3, 4),
                               chain_offset=1000, flags=flags)
        kw = dict(inv_mass=i["w"])
        if flags & _abi.FLAG_ACCUM_WELFORD:
            kw["welford"] = wf
        monkeypatch.setenv("MCMC_HMC_WS", "0")
        monkeypatch.delenv("MCMC_HMC_FORCE_CHUNKS", raising=False)
        ref = g.run(sampler, p, i["q0"], **kw)
        monkeypatch.setenv("MCMC_HMC_WS", "1")
        for nchunks in (None, "3", "9"):
            if nchunks:
                monkeypatch.setenv("MCMC_HMC_FORCE_CHUNKS", nchunks)
            got = g.run(sampler, p, i["q0"], **kw)
            for k in ref:
                assert np.array_equal(ref[k], got[k], equal_nan=True), (flags, nchunks, k)
    monkeypatch.delenv("MCMC_HMC_FORCE_CHUNKS", raising=False)
    monkeypatch.delenv("MCMC_HMC_WS", raising=False)
