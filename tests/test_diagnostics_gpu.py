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
