"""CPU-side checks of the C-ABI library: it loads, exports every declared symbol, the ctypes
mirror matches the header, and compute entries refuse to run without a GPU (no CPU fallback)."""
import ctypes as C
import os
import re
import subprocess

import numpy as np
import pytest

from mcmc_b200 import _abi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "mcmc_b200.h")


@pytest.mark.parametrize("variant", ["fast", "strict"])
def test_library_loads_and_exports_every_declared_symbol(variant):
    L = _abi.lib(variant)
    text = open(HEADER).read()
    declared = sorted(set(re.findall(r"\b(mcmc_[a-z0-9_]+)\s*\(", text)))
    assert declared == sorted(_abi.EXPORTS)
    for name in declared:
        assert getattr(L, name) is not None
    assert L.mcmc_abi_version() == _abi.ABI_VERSION


def test_struct_layout_matches_header(tmp_path):
    src = tmp_path / "sz.c"
    src.write_text('#include <stdio.h>\n#include <stddef.h>\n#include "%s"\nint main(){printf("%%zu %%zu %%zu %%zu %%zu\\n",'
                   'sizeof(McmcParams),sizeof(McmcBuffers),offsetof(McmcParams,step_size),'
                   'offsetof(McmcBuffers,welford_count),offsetof(McmcBuffers,workspace_bytes));return 0;}\n' % HEADER)
    exe = tmp_path / "sz"
    subprocess.check_call(["/usr/bin/gcc", str(src), "-o", str(exe)])
    got = [int(x) for x in subprocess.check_output([str(exe)]).split()]
    want = [C.sizeof(_abi.McmcParams), C.sizeof(_abi.McmcBuffers), _abi.McmcParams.step_size.offset,
            _abi.McmcBuffers.welford_count.offset, _abi.McmcBuffers.workspace_bytes.offset]
    assert got == want


def test_argument_validation_and_no_cpu_fallback():
    from mcmc_b200.hostrun import make_params, run_host
    L = _abi.lib("fast")
    q = np.zeros((2, 3))
    with pytest.raises(ValueError):  # unknown target -> the reference's ValueError (targets.py:527)
        run_host(_abi.SAMPLER_HMC, make_params(99, [], 2, 3, 1, num_steps=1, step_size=0.1), q)
    with pytest.raises(ValueError):  # unknown schedule -> KeyError in the reference (GRAHMC.py:122)
        run_host(_abi.SAMPLER_GRAHMC, make_params(0, [1.0], 2, 3, 1, num_steps=1, step_size=0.1, schedule_id=7), q)
    with pytest.raises(ValueError):  # injected mode without the arrays
        run_host(_abi.SAMPLER_HMC, make_params(0, [1.0], 2, 3, 1, num_steps=1, step_size=0.1,
                                               rng_mode=_abi.RNG_INJECTED), q)
    if L.mcmc_device_count() == 0:
        with pytest.raises(_abi.McmcError) as e:
            run_host(_abi.SAMPLER_HMC, make_params(0, [1.0], 2, 3, 1, num_steps=1, step_size=0.1), q)
        assert e.value.code == _abi.ECUDA and "no CPU fallback" in str(e.value)


def test_missing_library_fails_loudly(monkeypatch):
    monkeypatch.setitem(_abi.LIB_PATHS, "fast", "/nonexistent/libmcmc_b200.so")
    monkeypatch.setattr(_abi, "_LIBS", {})
    with pytest.raises(ImportError):
        _abi.lib("fast")


def test_product_package_never_touches_the_oracle():
    pkg = os.path.join(ROOT, "mcmc_b200")
    for dp, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                text = open(os.path.join(dp, f)).read()
                assert "import oracle" not in text and "from oracle" not in text and "liboracle" not in text, f
