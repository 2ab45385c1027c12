"""Measures the register-resident DFMA throughput of the GPU (the fp64 roofline denominator;
MEASURED_PEAKS.json has no fp64 entry).  Prints one JSON line."""
import ctypes as C
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from mcmc_b200 import _abi  # noqa: E402


def probe(iters=20000, reps=5):
    L = _abi.lib("fast")
    best = 0.0
    for _ in range(reps):
        fl, ms = C.c_double(0), C.c_float(0)
        _abi.check(L.mcmc_fp64_peak_probe(iters, C.byref(fl), C.byref(ms), None))
        best = max(best, fl.value / (ms.value * 1e-3) / 1e12)
    return best


if __name__ == "__main__":
    print(json.dumps({"fp64_dfma_tflops": probe()}))
