"""Device-pointer entries of the C ABI called the way a foreign binding would (raw pointers from torch storage): pointer
alignment, the MCMC_NO_SPIN schedule switch and the device handling of mcmc_run_host."""
import ctypes as C

import numpy as np
import pytest
import torch

from helpers import TARGET_CASES, GpuBackend, tid
from mcmc_b200 import _abi
from mcmc_b200.hostrun import make_params

pytestmark = pytest.mark.gpu


def _rwmh_device(L, q0, S, B, samples_view, seed=(4, 2)):
    Cn, D = q0.shape
    p = make_params(tid("standard_normal"), [1.0], Cn, D, S, B, 0, 0.5, seed=seed)
    out = dict(log_probs=torch.empty((S, Cn), dtype=torch.float64, device="cuda"),
               accept_count=torch.empty((Cn,), dtype=torch.int32, device="cuda"),
               final_q=torch.empty((Cn, D), dtype=torch.float64, device="cuda"))
    b = _abi.McmcBuffers()
    b.init_q = C.c_void_p(q0.data_ptr())
    b.samples = C.c_void_p(samples_view.data_ptr())
    for k, v in out.items():
        setattr(b, k, C.c_void_p(v.data_ptr()))
    _abi.check(L.mcmc_set_device(0))
    _abi.check(L.mcmc_rwmh_run(C.byref(p), C.byref(b), C.c_void_p(torch.cuda.current_stream().cuda_stream)))
    torch.cuda.synchronize()
    return out


@pytest.mark.parametrize("D", [20, 10, 7])
def test_rwmh_samples_pointer_that_is_only_8_byte_aligned(gpu_lib, D):
    """cp.async.bulk needs 16-byte aligned global addresses; a caller-owned samples buffer that is only 8-byte aligned
    must take the plain-store path and give the same bits (the header promises element alignment is enough)."""
    Cn, S, B = 333, 5, 2
    q0 = torch.from_numpy(np.random.default_rng(1).standard_normal((Cn, D))).cuda()
    buf = torch.zeros((S * Cn * D + 2,), dtype=torch.float64, device="cuda")
    base = buf.data_ptr()
    off_al = 0 if base % 16 == 0 else 1
    aligned = buf[off_al:off_al + S * Cn * D]
    assert aligned.data_ptr() % 16 == 0
    ra = _rwmh_device(gpu_lib, q0, S, B, aligned)
    want = aligned.clone()
    buf.zero_()
    mis = buf[off_al + 1:off_al + 1 + S * Cn * D]
    assert mis.data_ptr() % 16 == 8
    rm = _rwmh_device(gpu_lib, q0, S, B, mis)
    assert torch.equal(mis, want)
    for k in ra:
        assert torch.equal(ra[k], rm[k]), k
    assert torch.equal(want.reshape(S, Cn, D)[-1], ra["final_q"])


@pytest.mark.parametrize("sampler,L", [(_abi.SAMPLER_HMC, 6), (_abi.SAMPLER_GRAHMC, 6), (_abi.SAMPLER_NUTS, 0)])
def test_no_spin_switch_gives_the_same_bits(gpu_lib, oracle, monkeypatch, sampler, L):
    """MCMC_NO_SPIN=1 turns the chunked work queues off (no in-kernel wait on another block): a schedule change only."""
    name, D, Cn, S, B = "neals_funnel", 20, 40000, 6, 2      # more chains than resident threads of the NUTS grid
    q0 = TARGET_CASES[name][2](np.random.default_rng(9), Cn, D)
    p = oracle.make_params(tid(name), [], Cn, D, S, B, L, 0.05, gamma=1.0, steepness=0.5, schedule_id=1,
                           rng_mode=_abi.RNG_PHILOX, seed=(8, 9), max_tree_depth=5)
    g = GpuBackend("fast")
    monkeypatch.delenv("MCMC_NO_SPIN", raising=False)
    monkeypatch.setenv("MCMC_HMC_FORCE_CHUNKS", "3")
    ref = g.run(sampler, p, q0)
    monkeypatch.delenv("MCMC_HMC_FORCE_CHUNKS", raising=False)
    monkeypatch.setenv("MCMC_NO_SPIN", "1")
    got = g.run(sampler, p, q0)
    for k in ref:
        assert np.array_equal(ref[k], got[k], equal_nan=True), k


def test_run_host_sets_and_restores_the_device(gpu_lib):
    """mcmc_run_host makes `device` current for the call whatever the calling thread had, and puts the previous device
    back.  With one GPU the observable part is that a bad ordinal fails cleanly; with two, results match per device."""
    from mcmc_b200.hostrun import run_host
    q0 = np.random.default_rng(2).standard_normal((64, 20))
    p = make_params(tid("standard_normal"), [1.0], 64, 20, 4, 1, 4, 0.2, seed=(1, 1))
    r0 = run_host(_abi.SAMPLER_HMC, p, q0, device=0)
    with pytest.raises(_abi.McmcError) as e:
        run_host(_abi.SAMPLER_HMC, p, q0, device=gpu_lib.mcmc_device_count())
    assert e.value.code == _abi.ECUDA
    if gpu_lib.mcmc_device_count() >= 2:
        import threading
        res = {}

        def work(dev):
            res[dev] = run_host(_abi.SAMPLER_HMC, p, q0, device=dev)     # fresh thread: current device is 0
        ts = [threading.Thread(target=work, args=(d,)) for d in (1, 0)]
        [t.start() for t in ts]
        [t.join() for t in ts]
        for d in (0, 1):
            assert np.array_equal(res[d]["samples"], r0["samples"])
    _abi.check(gpu_lib.mcmc_host_arena_release())
    r1 = run_host(_abi.SAMPLER_HMC, p, q0, device=0)
    assert np.array_equal(r1["samples"], r0["samples"])
