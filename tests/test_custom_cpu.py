"""The extension point for user-supplied densities on a machine without a GPU: nvcc builds the plugin, the library
registers it, bad sources fail with the compiler's message, and compute still refuses to run without a device."""
import numpy as np
import pytest

from custom_cases import BROKEN, SCALED_NORMAL
from mcmc_b200 import _abi, custom


@pytest.fixture(scope="module")
def cache(tmp_path_factory):
    import os
    d = tmp_path_factory.mktemp("user_targets")
    old = os.environ.get("MCMC_B200_CACHE")
    os.environ["MCMC_B200_CACHE"] = str(d)
    yield d
    if old is None:
        os.environ.pop("MCMC_B200_CACHE", None)
    else:
        os.environ["MCMC_B200_CACHE"] = old


def test_compile_and_register_a_user_target(cache):
    t = custom.compile_target("scaled_normal", SCALED_NORMAL, dim=20, params=[2.0])
    tag = t.log_prob_fn
    assert tag.target_id >= 9 and tag.target_params == [2.0] and t.dim == 20 and t.true_mean is None
    again = custom.compile_target("scaled_normal", SCALED_NORMAL, dim=20, params=[3.0])     # same source: same plugin
    assert again.log_prob_fn.target_id == tag.target_id and again.log_prob_fn.target_params == [3.0]
    from mcmc_b200.benchmarks.targets import resolve_log_prob
    assert resolve_log_prob(t) is tag
    if _abi.lib("fast").mcmc_device_count() == 0:
        from mcmc_b200.hostrun import make_params, run_host
        with pytest.raises(_abi.McmcError) as e:     # registered, but there is still no CPU fallback
            run_host(_abi.SAMPLER_HMC, make_params(tag.target_id, [2.0], 2, 20, 1, num_steps=1, step_size=0.1), np.zeros((2, 20)))
        assert e.value.code == _abi.ECUDA
    from mcmc_b200.hostrun import make_params, run_host
    with pytest.raises(ValueError):                  # an id nobody registered
        run_host(_abi.SAMPLER_HMC, make_params(16, [], 2, 20, 1, num_steps=1, step_size=0.1), np.zeros((2, 20)))


def test_bad_sources_fail_loudly(cache):
    with pytest.raises(ValueError):
        custom.compile_target("nothing", "__device__ double f(double x) { return x; }", dim=3)
    with pytest.raises(RuntimeError) as e:
        custom.compile_target("broken", BROKEN, dim=3)
    assert "this_function_does_not_exist" in str(e.value)
    with pytest.raises(ValueError):
        custom.compile_target("too_many", SCALED_NORMAL, dim=3, params=list(range(9)))
