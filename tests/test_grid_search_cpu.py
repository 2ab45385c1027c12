"""Selection rule of run_trajectory_length_grid_search (run_benchmarks.py:402-413 and the failure branch :264-400) with
the per-L benchmark replaced by canned results: argmax of ESS per gradient among quality_pass runs, else among usable
runs, else an explicit failure record built from the least-bad run (lowest R-hat, then highest ESS)."""
import pytest

from mcmc_b200.benchmarks import grid_search as G
from mcmc_b200.benchmarks import targets as T


def _canned(table):
    def fake(sampler, target, target_name, key, n_chains, num_warmup, num_samples, schedule_type, num_steps, *a, **k):
        r = dict(table[num_steps])
        r.setdefault("num_steps", num_steps)
        r.setdefault("total_samples", num_samples)
        return r
    return fake


def _run(monkeypatch, table):
    monkeypatch.setattr(G, "run_single_benchmark_with_L", _canned(table))
    return G.run_trajectory_length_grid_search("hmc", T.standard_normal(10), "standard_normal", 0, 4, 100, 50, None,
                                               sorted(table), verbose=False)


def test_quality_pass_runs_win_over_more_efficient_usable_runs(monkeypatch):
    table = {8: dict(usable=True, quality_pass=False, ess_bulk_min=900.0, n_gradients=1000, rhat_max=1.03),
             16: dict(usable=True, quality_pass=True, ess_bulk_min=500.0, n_gradients=2000, rhat_max=1.005),
             32: dict(usable=True, quality_pass=True, ess_bulk_min=1200.0, n_gradients=4000, rhat_max=1.002),
             64: dict(usable=False, quality_pass=False, ess_bulk_min=5000.0, n_gradients=100, rhat_max=1.2)}
    best = _run(monkeypatch, table)
    assert best["num_steps"] == 32 and best["ess_per_gradient"] == pytest.approx(0.3)
    info = best["grid_search_info"]
    assert info["selected_L"] == 32 and info["selection_tier"] == "quality_pass" and info["has_usable"]
    assert [r["num_steps"] for r in info["all_results"]] == [8, 16, 32, 64]
    assert info["all_results"][0]["ess_per_gradient"] == pytest.approx(0.9)


def test_falls_back_to_usable_runs(monkeypatch):
    table = {8: dict(usable=True, quality_pass=False, ess_bulk_min=900.0, n_gradients=1000, rhat_max=1.03),
             16: dict(usable=True, quality_pass=False, ess_bulk_min=500.0, n_gradients=250, rhat_max=1.04),
             32: dict(error="boom", usable=False, quality_pass=False, ess_bulk_min=0.0)}
    best = _run(monkeypatch, table)
    assert best["num_steps"] == 16 and best["grid_search_info"]["selection_tier"] == "usable_only"
    rows = best["grid_search_info"]["all_results"]
    assert rows[2]["error"] == "boom" and rows[2]["ess_per_gradient"] == 0 and rows[2]["n_gradients"] == 0


def test_no_usable_run_returns_an_explicit_failure_from_the_least_bad_run(monkeypatch):
    table = {8: dict(usable=False, quality_pass=False, ess_bulk_min=300.0, n_gradients=100, rhat_max=1.30, step_size=0.1),
             16: dict(usable=False, quality_pass=False, ess_bulk_min=200.0, n_gradients=100, rhat_max=1.10, step_size=0.2),
             32: dict(usable=False, quality_pass=False, ess_bulk_min=900.0, n_gradients=100, rhat_max=1.10, step_size=0.3),
             64: dict(error="x", usable=False, quality_pass=False, ess_bulk_min=0.0)}
    f = _run(monkeypatch, table)
    assert f["grid_search_failed"] and not f["usable"] and f["error"] == "No trajectory length produced usable samples"
    assert f["num_steps"] == 32 and f["step_size"] == 0.3          # R-hat tie at 1.10 broken by the higher ESS
    assert f["rhat_max"] == 1.10 and f["ess_bulk_min"] == 900.0
    info = f["grid_search_info"]
    assert info["selected_L"] is None and info["least_bad_L"] == 32 and not info["has_usable"] and len(info["all_results"]) == 4


def test_n_gradients_defaults_to_samples_times_L(monkeypatch):
    best = _run(monkeypatch, {4: dict(usable=True, quality_pass=True, ess_bulk_min=100.0, rhat_max=1.0)})
    assert best["n_gradients"] == 50 * 4 and best["ess_per_gradient"] == pytest.approx(0.5)


def test_track_convergence_is_refused():
    with pytest.raises(NotImplementedError):
        G.run_single_benchmark_with_L("hmc", T.standard_normal(10), "standard_normal", 0, 4, 100, 50, None, 8,
                                      track_convergence=True)
