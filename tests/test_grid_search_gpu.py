"""run_trajectory_length_grid_search on the CUDA samplers (the reference's flow, run_benchmarks.py:198-935): per L a
2500-step adaptive warmup, one sampling launch, device diagnostics, gates, Sliced-W2."""
import pytest

pytestmark = pytest.mark.gpu


def test_grid_search_hmc_ill_conditioned_gaussian(gpu_lib):
    from mcmc_b200.benchmarks import targets as T
    from mcmc_b200.benchmarks.grid_search import run_trajectory_length_grid_search
    tgt = T.ill_conditioned_gaussian(20, condition_number=100.0)
    best = run_trajectory_length_grid_search("hmc", tgt, "ill_conditioned_gaussian", 3, 1024, 2500, 600, None,
                                             [4, 8, 16], verbose=False)
    info = best["grid_search_info"]
    assert info["has_usable"] and info["selected_L"] in (4, 8, 16), info
    rows = {r["num_steps"]: r for r in info["all_results"]}
    assert all(r["error"] is None for r in rows.values())
    for L, r in rows.items():
        assert r["n_gradients"] == 600 * L * 1024
        assert r["ess_per_gradient"] == pytest.approx(r["ess_bulk_min"] / r["n_gradients"])
        assert 0.3 < r["accept_rate"] < 0.95 and r["sample_time"] > 0 and r["min_ess_per_sec"] > 0
    pool = [r for r in rows.values() if r["quality_pass"]] or [r for r in rows.values() if r["usable"]]
    assert best["num_steps"] == max(pool, key=lambda r: r["ess_per_gradient"])["num_steps"]
    assert best["usable"] and best["rhat_max"] < 1.05 and best["ess_bulk_min"] >= 400 and best["sliced_w2"] is not None
    assert best["mass_matrix_learned"] and 0.5 < best["mass_matrix_min"] < 2.0 and best["mass_matrix_max"] > 50.0
    assert best["z_score_threshold"] == pytest.approx(3.0233, abs=1e-3)


@pytest.mark.parametrize("sampler,sched", [("nuts", None), ("grahmc", "tanh"), ("rwmh", None)])
def test_single_benchmark_other_samplers(gpu_lib, sampler, sched):
    from mcmc_b200.benchmarks import targets as T
    from mcmc_b200.benchmarks.grid_search import run_single_benchmark_with_L
    tgt = T.standard_normal(10)
    S = 4000 if sampler == "rwmh" else 400
    r = run_single_benchmark_with_L(sampler, tgt, "standard_normal", 5, 512, 2500, S, sched, 8, verbose=False)
    assert r.get("error") is None, r.get("error")
    assert r["usable"], (r["rhat_max"], r["ess_bulk_min"], r["ess_tail_min"])
    if sampler == "nuts":
        assert r["num_steps"] is None and r["avg_tree_depth"] >= 1.0 and r["n_gradients"] >= S * 512
    elif sampler == "rwmh":
        assert r["n_gradients"] == 0 and r["ess_per_gradient"] == 0 and 0.1 < r["accept_rate"] < 0.45
    else:
        assert r["schedule"] == "tanh" and r["n_gradients"] == S * 8 * 512 and "gamma" in r
