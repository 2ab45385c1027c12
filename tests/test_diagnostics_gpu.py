"""Device diagnostics (mcmc_b200/csrc/diag.cu through the C ABI) against the NumPy restatement of ArviZ."""
import numpy as np
import pytest
import torch

from oracle import diagnostics_np as O

pytestmark = pytest.mark.gpu


def _mcmc_like(rng, S, C, D, rho=0.6, p_stay=0.25):
    """AR(1) chains with repeated values (rejections) so that ties exist, chain-specific offsets."""
    x = np.zeros((S, C, D))
    x[0] = rng.standard_normal((C, D))
    for s in range(1, S):
        prop = rho * x[s - 1] + np.sqrt(1 - rho ** 2) * rng.standard_normal((C, D))
        stay = rng.random((C, 1)) < p_stay
        x[s] = np.where(stay, x[s - 1], prop)
    return x + 0.05 * rng.standard_normal((1, C, D))


def test_rank_normalize_with_ties(gpu_lib):
    from mcmc_b200 import diagnostics as DG
    rng = np.random.default_rng(0)
    x = np.round(rng.standard_normal((64, 37)), 1)          # heavy ties
    z, vals = DG.rank_normalize(torch.from_numpy(x).cuda())
    np.testing.assert_allclose(z.cpu().numpy(), O.z_scale(x), rtol=0, atol=2e-13)
    assert np.array_equal(vals.cpu().numpy(), np.sort(x.reshape(-1)))


def test_rank_normalize_constant_and_stuck_columns(gpu_lib):
    """A column that never moved is ONE tie run of n elements: the run boundaries come from binary searches, so 4 M equal
    values cost the same as 4 M distinct ones (a linear scan per element would be O(n^2) and effectively hang)."""
    from mcmc_b200 import diagnostics as DG
    n = 1 << 22
    z, _ = DG.rank_normalize(torch.full((n,), 3.25, dtype=torch.float64, device="cuda"))
    torch.cuda.synchronize()
    assert float(z.abs().max()) < 1e-6                      # every element gets the middle rank -> Phi^-1(~0.5) ~ 0
    assert torch.equal(z, z[0].expand_as(z))
    # half the column stuck at one value, the rest distinct: against the NumPy restatement
    rng = np.random.default_rng(3)
    x = rng.standard_normal(4096)
    x[::2] = 0.125
    z, _ = DG.rank_normalize(torch.from_numpy(x).cuda())
    np.testing.assert_allclose(z.cpu().numpy(), O.z_scale(x), rtol=0, atol=2e-13)


def _fold_cases():
    rng = np.random.default_rng(5)
    yield "normal_100003", rng.standard_normal(100003)
    for n in (1, 2, 3, 4, 5, 6):
        yield f"tiny_{n}", rng.standard_normal(n)
    yield "heavy_ties_digits", rng.integers(0, 10, 50000).astype(np.float64)          # ties on both sides and across the fold
    yield "rounded_normal_2d", np.round(rng.standard_normal((64, 37)), 1)
    yield "constant", np.full(4097, 3.25)
    yield "symmetric_exact", np.concatenate([np.arange(-300, 301) / 8.0, np.arange(-300, 301) / 8.0])   # every distance twice per side
    yield "median_repeated", np.concatenate([rng.standard_normal(999), np.zeros(500), -rng.random(300), rng.random(300)])
    yield "signed_zeros", np.array([0.0, -0.0, 1.0, -1.0, 0.0, -0.0, 2.0, -2.0, -0.0])
    yield "two_values", np.array([1.0] * 700 + [2.0] * 301)
    yield "one_sided", np.concatenate([np.zeros(600), rng.random(400)])                   # median = minimum: empty left side
    yield "wide_range", rng.standard_normal(20001) * 10.0 ** rng.integers(-8, 8, 20001)
    yield "stuck_chains", _mcmc_like(rng, 128, 64, 1, p_stay=0.6)[:, :, 0]


@pytest.mark.parametrize("name", [n for n, _ in _fold_cases()])
def test_rank_normalize_fold_is_bit_identical_to_two_sorts(gpu_lib, name):
    """mcmc_rank_normalize_fold (one sort; the folded ranks from the two monotone sides of the median) against the
    straightforward path — sort, rank, fold around the median, sort again, rank — and against the NumPy restatement of
    ArviZ.  Ranks are exact half-integers, so the z-scores must agree bit for bit."""
    from mcmc_b200 import diagnostics as DG
    x = dict(_fold_cases())[name]
    xt = torch.from_numpy(np.ascontiguousarray(x)).cuda()
    zb, zf, vals = DG.rank_normalize_fold(xt)
    z0, v0 = DG.rank_normalize(xt)
    n = v0.numel()
    med = 0.5 * (v0[(n - 1) // 2] + v0[n // 2])
    zf0, _ = DG.rank_normalize((xt - med).abs())
    assert torch.equal(vals, v0)
    assert torch.equal(zb, z0), name
    assert torch.equal(zf, zf0), name
    assert zb.shape == xt.shape and zf.shape == xt.shape
    if x.size >= 4:
        xa = np.asarray(x, dtype=np.float64)
        np.testing.assert_allclose(zf.cpu().numpy(), O.z_scale(np.abs(xa - np.median(xa))), rtol=0, atol=2e-13)


def test_rank_normalize_fold_large_and_errors(gpu_lib):
    import ctypes as C
    from mcmc_b200 import _abi, diagnostics as DG
    g = torch.Generator(device="cuda").manual_seed(11)
    x = torch.randn((1 << 22) + 7, dtype=torch.float64, device="cuda", generator=g)
    m = x.numel() // 5
    x[0:5 * m:5] = x[1:5 * m:5]                                                           # 20% exact duplicates
    zb, zf, vals = DG.rank_normalize_fold(x)
    z0, v0 = DG.rank_normalize(x)
    med = 0.5 * (v0[(x.numel() - 1) // 2] + v0[x.numel() // 2])
    assert torch.equal(zb, z0) and torch.equal(zf, DG.rank_normalize((x - med).abs())[0]) and torch.equal(vals, v0)
    L = _abi.lib("fast")
    nb = C.c_size_t(0)
    assert L.mcmc_rank_normalize_fold_workspace(1 << 31, C.byref(nb)) == _abi.ENOTSUP
    assert L.mcmc_rank_normalize_fold_workspace(1000, C.byref(nb)) == 0 and nb.value > 8000
    p = lambda t: C.c_void_p(t.data_ptr())
    small = torch.empty(16, dtype=torch.uint8, device="cuda")
    assert L.mcmc_rank_normalize_fold(p(x), 1000, p(vals), p(zb), p(zf), p(small), 16, None) == _abi.EWORKSPACE
    assert L.mcmc_rank_normalize_fold(None, 1000, p(vals), p(zb), p(zf), p(small), 16, None) == _abi.EINVAL
    assert L.mcmc_rank_normalize_fold(None, 0, None, None, None, None, 0, None) == 0


@pytest.mark.parametrize("n,C", [(50, 7), (128, 100), (257, 33), (900, 40)])
def test_split_chain_stats_vs_numpy(gpu_lib, n, C):
    """n=900 exceeds the shared-memory tile and takes the global-scratch variant of the kernel."""
    from mcmc_b200 import diagnostics as DG
    rng = np.random.default_rng(n)
    y = _mcmc_like(rng, 2 * n, C, 1)[:, :, 0] * 3.0 + 10.0
    acov, stats = DG.split_chain_stats(torch.from_numpy(y).cuda(), n, n)
    seqs = np.vstack([y[:n].T, y[n:].T])                    # (2C, n)
    want = O.autocov(seqs).sum(axis=0)
    np.testing.assert_allclose(acov.cpu().numpy(), want, rtol=1e-10, atol=1e-10 * abs(want[0]))
    m = seqs.mean(axis=1)
    np.testing.assert_allclose(stats.cpu().numpy(), [2 * C, m.mean(), ((m - m.mean()) ** 2).sum()], rtol=1e-11)
    # strided input (a column of a (S, C, D) array) and a lag cap
    full = np.stack([y, -y, 2 * y], axis=2)
    a2, s2 = DG.split_chain_stats(torch.from_numpy(full).cuda()[:, :, 0], n, n, 5)
    np.testing.assert_array_equal(a2.cpu().numpy(), acov.cpu().numpy()[:5])
    np.testing.assert_array_equal(s2.cpu().numpy(), stats.cpu().numpy())


@pytest.mark.parametrize("S,C,D", [(100, 8, 3), (101, 4, 2), (64, 300, 2)])
def test_compute_diagnostics_vs_oracle(gpu_lib, S, C, D):
    from mcmc_b200 import diagnostics as DG
    x = _mcmc_like(np.random.default_rng(S + C), S, C, D)
    got = DG.compute_diagnostics(torch.from_numpy(x).cuda())
    want = O.compute_diagnostics(x)
    for k in ("rhat", "ess_bulk", "ess_tail"):
        np.testing.assert_allclose(got[k], want[k], rtol=1e-9, err_msg=k)
    for k in ("mean", "sd", "mcse_mean"):
        np.testing.assert_allclose(got["summary"][k], want["summary"][k], rtol=1e-9, atol=1e-12, err_msg=k)
    for k in ("rhat_max", "rhat_mean", "ess_bulk_min", "ess_bulk_mean", "ess_tail_min", "ess_tail_mean"):
        assert got[k] == pytest.approx(want[k], rel=1e-9)
    np.testing.assert_allclose(got["mean_estimate"], want["mean_estimate"], atol=1e-12)
    np.testing.assert_allclose(got["std_estimate"], want["std_estimate"], rtol=1e-10)


def test_diagnostics_reject_cpu_tensors(gpu_lib):
    from mcmc_b200 import diagnostics as DG
    with pytest.raises(RuntimeError):
        DG.compute_diagnostics(torch.zeros(10, 2, 2, dtype=torch.float64))


def test_sliced_w2_vs_numpy(gpu_lib):
    from mcmc_b200.benchmarks import metrics as M
    from mcmc_b200.random import PRNGKey, fold_in, normal
    rng = np.random.default_rng(5)
    a = rng.standard_normal((4000, 6))
    b = rng.standard_normal((4000, 6)) * 1.3 + 0.2
    key = PRNGKey(30)
    dirs = np.concatenate([normal(fold_in(key, i), (100, 6)) for i in range(3)])[:250]
    dirs /= np.linalg.norm(dirs, axis=1, keepdims=True)
    got = M.sliced_wasserstein_distance(a, b, n_projections=250, key=key)
    assert got == pytest.approx(O.sliced_wasserstein_np(a, b, dirs), rel=1e-10)
    # unequal sample counts go through quantile matching (metrics.py:66-72)
    got2 = M.sliced_wasserstein_distance(a, b[:3000], n_projections=100, key=key)
    assert got2 == pytest.approx(O.sliced_wasserstein_np(a, b[:3000], dirs[:100]), rel=1e-9)
    assert M.compute_sliced_w2(torch.from_numpy(a).cuda(), "rosenbrock", 6) is None
    w = M.compute_sliced_w2(torch.from_numpy(a.reshape(40, 100, 6)).cuda(), "standard_normal", 6, n_reference=4000)
    assert 0.0 < w < 0.08
