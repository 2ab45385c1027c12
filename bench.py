#!/usr/bin/env python
"""bench.py — fp64 grad-evals/sec of the sampling hot path (BASELINE.json metric).

Workload at every N (BASELINE.json configs[1]): HMC, diagonal mass, 20-D rosenbrock,
65,536 chains PER GPU (weak scaling: rank r owns global chains [r*C, (r+1)*C), Philox streams
keyed on the global chain id), L grid 16/32/64/128.  One "step" = one pass of the hot path over
the batch: four hmc_run calls (one per L) with S stored iterations each, i.e. C*S*(16+32+64+128)
leapfrog gradient evaluations (grad-eval = one leapfrog step of one chain,
run_benchmarks.py:733 of the reference).

  value  device-resident inputs, CUDA events on the launching stream, max over ranks
  e2e    the public API (mcmc_b200.samplers.hmc_run) called with a pinned HOST init_position;
         the timed region contains the H2D copy and the D2H read of EVERYTHING the reference's
         hmc_run returns (samples, log_probs, accept_rate, final state) into pinned host memory
  --impl reference   the reference's JAX path cannot run here (jax is not installed and cannot be:
         no wheel, no network), so this arm times the CPU restatement of the reference
         (oracle/mcmc_oracle.c, OpenMP over chains, all host threads) on a bounded sample of the
         same workload.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402

D = 20
CHAINS_PER_GPU = 65536
L_GRID = (16, 32, 64, 128)
S_STORED = 64            # stored iterations per hmc_run call (705 MB of samples per call)
STEP_SIZE = 0.02         # fixed epsilon, see DESIGN.md (short dual-averaging run, accept ~0.8)
INV_MASS = np.linspace(0.5, 2.0, D)  # a non-trivial diagonal M^-1 to exercise the path
SEED = (0, 0)
# algorithmic fp64 flops of THIS kernel's formulation (FMA = 2), DESIGN.md "flop model"
# dram__bytes_read.sum + dram__bytes_write.sum of one hmc_kernel launch at the bench shape (ncu --set full captures of
# L=16 and L=128 of round 2: 804.3 MB and 1049.3 MB, profiles/ncu_hmc_L{16,128}_r02c.txt; the L=128 launch is the 168-register
# instantiation, whose register spills add ~290 MB of local-memory write-back) — a constant from the profile, not a live
# measurement; the algorithmic bytes are the stored samples + log-probs, the initial state read and the final q / g written
NCU_TRAFFIC_BYTES_PER_LAUNCH = 926.8e6
NCU_TRAFFIC_SOURCE = ("profiles/ncu_hmc_L16_r02c.txt (804.3 MB), profiles/ncu_hmc_L128_r02c.txt (1049.3 MB incl. spill write-back); "
                      "mean of the two launches")
F_STEP = 231.0           # per leapfrog step of one 20-D rosenbrock chain: 80 integrator (2 DFMA/dim) + 151 gradient (75 DFMA + 1 DMUL)
F_ITER = 190.0           # per MH iteration outside the trajectory (momentum scaling, two kinetic energies, value)


def init_positions(rank: int) -> np.ndarray:
    rng = np.random.default_rng(1234 + rank)
    return 1.0 + 0.5 * rng.standard_normal((CHAINS_PER_GPU, D))  # targets.py:393 init_sampler


class ClockSampler(threading.Thread):
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region."""

    def __init__(self, index: int):
        super().__init__(daemon=True)
        self.index = index
        self.period = float(os.environ.get("MCMC_BENCH_CLOCK_PERIOD", "0.01"))
        self.rows = []
        self._stop_evt = threading.Event()
        self._armed = threading.Event()   # rows are kept only while armed (= during the timed region)
        self.ready = threading.Event()    # NVML initialised: nvmlInit can hold driver locks for milliseconds and
                                          # must not run concurrently with the first timed launches

    def run(self):
        try:
            import pynvml
            pynvml.nvmlInit()
            h = pynvml.nvmlDeviceGetHandleByIndex(self.index)
            mx = pynvml.nvmlDeviceGetMaxClockInfo(h, pynvml.NVML_CLOCK_SM)
            R = pynvml
            bits = [(R.nvmlClocksThrottleReasonHwSlowdown, "hw_slowdown"),
                    (R.nvmlClocksThrottleReasonHwThermalSlowdown, "hw_thermal_slowdown"),
                    (R.nvmlClocksThrottleReasonSwThermalSlowdown, "sw_thermal_slowdown"),
                    (R.nvmlClocksThrottleReasonSwPowerCap, "sw_power_cap")]
            pynvml.nvmlDeviceGetClockInfo(h, pynvml.NVML_CLOCK_SM)
            self.ready.set()
            while not self._stop_evt.is_set():
                if self._armed.is_set():
                    sm = pynvml.nvmlDeviceGetClockInfo(h, pynvml.NVML_CLOCK_SM)
                    rs = pynvml.nvmlDeviceGetCurrentClocksThrottleReasons(h)
                    self.rows.append([str(sm), str(mx)] + ["active" if rs & b else "not active" for b, _ in bits])
                self._stop_evt.wait(self.period)
            return
        except Exception:
            pass
        self.ready.set()
        q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
        while not self._stop_evt.is_set():
            if not self._armed.is_set():
                self._stop_evt.wait(0.01)
                continue
            try:
                o = subprocess.run(["nvidia-smi", f"--id={self.index}", f"--query-gpu={q}", "--format=csv,noheader,nounits"],
                                   capture_output=True, text=True, timeout=5).stdout.strip()
                if o:
                    self.rows.append([x.strip() for x in o.split(",")])
            except Exception:
                pass
            self._stop_evt.wait(0.1)

    def arm(self):
        self._armed.set()

    def stop(self):
        self._stop_evt.set()
        self.join(timeout=6)
        sm = [float(r[0]) for r in self.rows if r and r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if len(r) > 1 and r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({names[i] for r in self.rows if len(r) >= 6 for i in range(4) if r[2 + i].lower() == "active"})
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(sm)}


# ----------------------------------------------------------------------------------------------
# CPU arm: the oracle port of the reference (bench.py is one of the few places allowed to run it)

_NATIVE = {}
_JSON_FD = {"fd": None}     # the real stdout when fd 1 has been pointed at stderr (N > 1: NCCL logs to stdout)


def emit_json(line: dict) -> None:
    text = json.dumps(line) + "\n"
    if _JSON_FD["fd"] is None:
        sys.stdout.write(text)
        sys.stdout.flush()
    else:
        os.write(_JSON_FD["fd"], text.encode())


def native_oracle():
    """oracle/mcmc_oracle.c compiled on THIS host with -O3 -march=native and FMA contraction (the checker build is -O2,
    generic x86-64, -ffp-contract=off): the CPU arm is timed on the fast build."""
    from oracle import oracle as O
    if "so" not in _NATIVE:
        import tempfile
        path = os.path.join(tempfile.mkdtemp(prefix="mcmc_oracle_native_"), "liboracle_native.so")
        try:
            O.build_native(path)
            O.use_library(path)
            _NATIVE["so"] = path
            _NATIVE["flags"] = "-O3 -march=native -ffp-contract=fast -fopenmp"
        except Exception as e:   # no compiler on the box: time the checker build and say so
            _NATIVE["so"] = None
            _NATIVE["flags"] = f"-O2 -ffp-contract=off -fopenmp (native build failed: {type(e).__name__})"
        # torchrun exports OMP_NUM_THREADS=1 to every worker: the CPU arm sizes its OpenMP team to the host's cores itself
        _NATIVE["cores"] = O.use_all_cores()
    return O


def cpu_pass(chains: int, samples: int, threads_note: bool = True) -> dict:
    """One pass over the L grid on `chains` chains with the C oracle (all OpenMP threads)."""
    from mcmc_b200 import _abi
    O = native_oracle()
    q0 = init_positions(0)[:chains]
    t0 = time.perf_counter()
    n_grad = 0
    for L in L_GRID:
        p = O.make_params(_abi.TARGET_IDS["rosenbrock"], [0.1], chains, D, samples, 0, L, STEP_SIZE,
                          rng_mode=_abi.RNG_PHILOX, seed=SEED)
        r = O.run(_abi.SAMPLER_HMC, p, q0, inv_mass=INV_MASS)
        n_grad += int(r["counters"][1])
    dt = time.perf_counter() - t0
    return {"seconds": dt, "grad_evals": n_grad, "value": n_grad / dt, "cores": O.num_threads(), "flags": _NATIVE["flags"]}


def ensure_built() -> None:
    """The harness (not the product path, which fails loudly) builds the libraries when a fresh checkout has none.
    Local rank 0 builds; the other ranks wait for the digest stamps the build writes after linking."""
    from mcmc_b200 import _abi
    csrc = os.path.join(ROOT, "mcmc_b200", "csrc")
    need = list(_abi.LIB_PATHS.values()) + [os.path.join(ROOT, "oracle", "liboracle.so")]
    stamps = [os.path.join(csrc, "build", v, "digest.txt") for v in ("fast", "strict")]
    if all(os.path.exists(p) for p in need):
        return
    if int(os.environ.get("LOCAL_RANK", "0")) == 0:
        import __graft_entry__ as g
        g.build()
    else:
        t0 = time.time()
        while not all(os.path.exists(p) for p in need + stamps) and time.time() - t0 < 1200:
            time.sleep(2.0)


def run_reference_arm(args) -> None:
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    chains, samples = 16384, 32   # bounded sample: 16,384 of the 65,536 chains, 32 of the 64 stored iterations (~0.5 s per step)
    for _ in range(args.warmup):
        cpu_pass(chains, samples)
    tot_t, tot_g, cores = 0.0, 0, 1
    for _ in range(args.steps):
        r = cpu_pass(chains, samples)
        tot_t += r["seconds"]
        tot_g += r["grad_evals"]
        cores = r["cores"]
    val = tot_g / tot_t
    sample = f"{chains} of {CHAINS_PER_GPU} chains x {samples} of {S_STORED} stored iterations x L grid {list(L_GRID)} per step"
    line = {
        "impl": "reference", "metric": "fp64_grad_evals_per_sec", "value": val, "unit": "grad-evals/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * tot_t / max(args.steps, 1),
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": workload_config(args.gpus),
        "cpu_baseline": {"value": val, "unit": "grad-evals/s", "cores": cores, "kind": "port", "sample": sample,
                         "note": "reference JAX CPU path unavailable on this box (jax not installed, no network); "
                                 "this is the C restatement oracle/mcmc_oracle.c, OpenMP over chains, gcc " + _NATIVE.get("flags", "")},
        "e2e": {"value": val, "unit": "grad-evals/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


def workload_config(n_gpus: int) -> dict:
    return {"workload": "BASELINE configs[1]: HMC diagonal-mass, 20-D rosenbrock(scale=0.1), 65,536 chains per GPU, "
                        "L grid 16/32/64/128, one step = 4 hmc_run calls (one per L)",
            "chains_per_gpu": CHAINS_PER_GPU, "n_dim": D, "L_grid": list(L_GRID), "stored_iterations_per_call": S_STORED,
            "step_size": STEP_SIZE, "rng": "philox4x32-10 in-kernel", "parallelism": f"chains sharded x{n_gpus}",
            "l2": "each call writes 705 MB of samples (> 126 MB L2); the kernel is fp64-pipe bound"}


# ----------------------------------------------------------------------------------------------
# GPU arm

def fp64_peak_tflops(reps: int = 3) -> float:
    """Register-resident DFMA probe (MEASURED_PEAKS.json has no fp64 entry): best of `reps`."""
    import ctypes as C
    from mcmc_b200 import _abi
    L = _abi.lib("fast")
    best = 0.0
    for _ in range(reps):
        fl, ms = C.c_double(0), C.c_float(0)
        _abi.check(L.mcmc_fp64_peak_probe(20000, C.byref(fl), C.byref(ms), None))
        best = max(best, fl.value / (ms.value * 1e-3) / 1e12)
    return best


def run_gpu_arm(args) -> None:
    import torch
    import torch.distributed as dist
    from mcmc_b200 import _abi
    from mcmc_b200.benchmarks.targets import rosenbrock
    from mcmc_b200.samplers import hmc_run

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise RuntimeError("bench.py needs a CUDA device (the product path has no CPU fallback)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        # NCCL's INFO log (communicator ranks, NVLS / ring setup) goes to stderr; stdout carries rank 0's ONE JSON line.
        # A value the launcher already exported is respected.
        # (the GPU boxes export NCCL_DEBUG=VERSION: only the banner; anything else a launcher set is respected)
        if os.environ.get("NCCL_DEBUG", "VERSION").upper() in ("", "VERSION"):
            os.environ["NCCL_DEBUG"] = "INFO"
            os.environ.setdefault("NCCL_DEBUG_SUBSYS", "INIT,ENV")
        # NCCL logs to the process's stdout (fd 1): keep a duplicate of the real stdout for the JSON line and point fd 1
        # at stderr for the rest of the run
        sys.stdout.flush()
        _JSON_FD["fd"] = os.dup(1)
        os.dup2(2, 1)
        dist.init_process_group("nccl", device_id=dev)
        from mcmc_b200 import dist as _mdist
        _mdist.warm_up(dev)          # NCCL's lazy channel set-up happens here, not inside a timed leg
    _abi.check(_abi.lib("fast").mcmc_set_device(local))

    target = rosenbrock(D, scale=0.1)
    q0_host = torch.from_numpy(init_positions(rank)).pin_memory()
    q0_dev = q0_host.to(dev)
    w_dev = torch.from_numpy(INV_MASS).to(dev)
    C = CHAINS_PER_GPU
    offset = rank * C

    def one_call(L, init, it):
        # a fresh Philox key per (step, L) so no two calls replay the same stream
        key = np.array([(SEED[0] + it) & 0xFFFFFFFF, (SEED[1] + L) & 0xFFFFFFFF], dtype=np.uint64).astype(np.uint32)
        return hmc_run(key, target.log_prob_fn, init, STEP_SIZE, L, S_STORED, burn_in=0, inv_mass_matrix=w_dev,
                       _chain_offset=offset)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)

    peak = fp64_peak_tflops()
    sampler = ClockSampler(local) if rank == 0 else None
    if sampler:
        sampler.start()
        sampler.ready.wait(timeout=30)

    # ---- device-resident loop (value)
    out = None
    for i in range(args.warmup):
        for L in L_GRID:
            # keep the previous result alive while the next call allocates, exactly like the timed loop: torch's
            # caching allocator then owns BOTH 705 MB sample blocks before timing starts (a cudaMalloc inside the
            # timed region stalls the stream for milliseconds)
            out = one_call(L, q0_dev, -1 - i)
    barrier()
    if sampler:
        sampler.arm()
    evs = []
    host_ms = {}
    acc_sum = 0.0
    t_start = torch.cuda.Event(enable_timing=True)
    t_end = torch.cuda.Event(enable_timing=True)
    t_start.record()
    for i in range(args.steps):
        for L in L_GRID:
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            h0 = time.perf_counter()
            out = one_call(L, q0_dev, i)
            host_ms[L] = host_ms.get(L, 0.0) + 1e3 * (time.perf_counter() - h0)
            e1.record()
            evs.append((L, e0, e1))
    t_end.record()
    barrier()
    dev_ms = t_start.elapsed_time(t_end)
    if os.environ.get("MCMC_BENCH_DEBUG"):
        print("host ms per call:", {L: round(v / args.steps, 3) for L, v in host_ms.items()}, file=sys.stderr)
    clocks = sampler.stop() if sampler else None
    acc_sum = float(out[2].mean().item())
    kern_ms = {L: 0.0 for L in L_GRID}
    for L, e0, e1 in evs:
        kern_ms[L] += e0.elapsed_time(e1)
    grad_per_step = C * S_STORED * sum(L_GRID)
    t = torch.tensor([dev_ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    dev_ms_max = float(t.item())
    value = world * grad_per_step * args.steps / (dev_ms_max * 1e-3)

    # ---- end-to-end loop: pinned host init in, every reference output back to pinned host memory
    host_out = {
        "samples": torch.empty((S_STORED, C, D), dtype=torch.float64).pin_memory(),
        "log_probs": torch.empty((S_STORED, C), dtype=torch.float64).pin_memory(),
        "accept_rate": torch.empty((C,), dtype=torch.float32).pin_memory(),
        "final_q": torch.empty((C, D), dtype=torch.float64).pin_memory(),
        "final_lp": torch.empty((C,), dtype=torch.float64).pin_memory(),
        "final_g": torch.empty((C, D), dtype=torch.float64).pin_memory(),
    }
    copy_stream = torch.cuda.Stream(dev)
    h2d = q0_host.numel() * 8
    d2h = sum(v.numel() * v.element_size() for v in host_out.values())

    def e2e_step(i):
        for L in L_GRID:
            smp, lps, ar, st = one_call(L, q0_host, 1000 + i)   # H2D of the pinned init happens inside
            ready = torch.cuda.Event()
            ready.record()
            copy_stream.wait_event(ready)
            with torch.cuda.stream(copy_stream):               # D2H overlaps the next L's kernel
                host_out["samples"].copy_(smp, non_blocking=True)
                host_out["log_probs"].copy_(lps, non_blocking=True)
                host_out["accept_rate"].copy_(ar, non_blocking=True)
                host_out["final_q"].copy_(st.position, non_blocking=True)
                host_out["final_lp"].copy_(st.log_prob, non_blocking=True)
                host_out["final_g"].copy_(st.grad_log_prob, non_blocking=True)
                for x in (smp, lps, ar, st.position, st.log_prob, st.grad_log_prob):
                    x.record_stream(copy_stream)

    e2e_steps = max(1, min(args.steps, 5))
    e2e_step(-1)
    barrier()
    w0 = time.perf_counter()
    for i in range(e2e_steps):
        e2e_step(i)
    barrier()
    e2e_s = time.perf_counter() - w0
    t = torch.tensor([e2e_s], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    e2e_value = world * grad_per_step * e2e_steps / float(t.item())
    assert np.isfinite(host_out["log_probs"].numpy()).all()

    # ---- second half of the metric: min-ESS/sec per L (run_benchmarks.py:179,188,712: ess_bulk_min / sample_time), from
    # one more call per L at the bench shape; the bulk ESS is over ALL chains of the job (distributed diagnostics at N>1)
    # and rhat_max is printed beside it: at eps = 0.02 and 64 draws these chains have NOT converged
    from mcmc_b200 import diagnostics as DG
    from mcmc_b200 import dist as mdist
    ess = {}
    for L in L_GRID:
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        o = one_call(L, q0_dev, 2000)
        e1.record()
        torch.cuda.synchronize(dev)
        tt = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        dg = DG.compute_diagnostics(o[0], distributed=world > 1)
        ess[L] = {"ess_bulk_min": dg["ess_bulk_min"], "rhat_max": dg["rhat_max"], "kernel_ms": float(tt.item()),
                  "min_ess_per_sec": dg["ess_bulk_min"] / (float(tt.item()) * 1e-3)}
        del o

    # ---- a communicating leg at every N, outside `value`: (1) one full run_adaptive_warmup of BASELINE configs[2]
    # (GRAHMC tanh on the 20-D funnel, 32,768 chains per GPU, 2500 steps = 26 dual-averaging batches, device Welford;
    # per batch the accept statistic and per window the chain variances cross the ranks), (2) R-hat / ESS of the last
    # call's samples over all ranks' chains
    from mcmc_b200.benchmarks.targets import neals_funnel
    from mcmc_b200.tuning import run_adaptive_warmup
    funnel = neals_funnel(D)
    cw = 32768
    qw = torch.as_tensor(np.asarray(funnel.init_sampler(77 + rank, cw)), dtype=torch.float64).to(dev)
    comm = {}
    for rep in range(2):                         # first pass warms every kernel variant and NCCL payload size
        n0 = dict(mdist.COUNTERS)
        barrier()
        w0 = time.perf_counter()
        eps_w, _, _, info_w = run_adaptive_warmup("grahmc", funnel.log_prob_fn, None, qw, key=5, num_warmup=2500,
                                                  schedule_type="tanh", num_steps=16, verbose=False, tune_friction=False,
                                                  chain_offset=rank * cw)
        barrier()
        tw = torch.tensor([time.perf_counter() - w0], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(tw, op=dist.ReduceOp.MAX)
        comm = {"warmup_ms": 1e3 * float(tw.item()), "warmup_chains_per_gpu": cw,
                "n_allreduce": mdist.COUNTERS["all_reduce"] - n0["all_reduce"], "da_updates": info_w["num_da_updates"],
                "final_step_size": eps_w,
                "warmup_grad_evals_per_sec": world * cw * 2500 * 16 / float(tw.item())}
    for rep in range(2):
        barrier()
        w0 = time.perf_counter()
        dgl = DG.compute_diagnostics(out[0], distributed=world > 1)
        barrier()
        td_ = torch.tensor([time.perf_counter() - w0], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(td_, op=dist.ReduceOp.MAX)
        comm["diagnostics_ms"] = 1e3 * float(td_.item())
    comm["diagnostics_shape"] = [S_STORED, world * C, D]
    comm["diagnostics_rhat_max"] = dgl["rhat_max"]

    if rank == 0:
        tot_kern_ms = sum(kern_ms.values())
        flops = sum(C * S_STORED * (L * F_STEP + F_ITER) for L in L_GRID) * args.steps
        achieved = flops / (tot_kern_ms * 1e-3) / 1e12
        wb_bytes = C * S_STORED * (D + 1) * 8
        line = {
            "metric": "fp64_grad_evals_per_sec", "value": value, "unit": "grad-evals/s", "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": dev_ms_max / args.steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": workload_config(world),
            "per_L": {str(L): {"ms_per_call": kern_ms[L] / args.steps,
                               "grad_evals_per_sec": C * S_STORED * L * args.steps / (kern_ms[L] * 1e-3)} for L in L_GRID},
            "roofline": {"bound": "fp64", "achieved": achieved, "peak": peak, "unit": "TFLOP/s", "frac": achieved / peak,
                         "traffic": NCU_TRAFFIC_BYTES_PER_LAUNCH, "traffic_source": NCU_TRAFFIC_SOURCE,
                         "algorithmic_bytes_per_launch": wb_bytes + 3 * C * D * 8 + 2 * C * 8,
                         "kernel": "mcmc::hmc_kernel<6, 20, false, 64, false, 1> at L=16/32/64 and <..., 6> (6 blocks/SM) at L=128 (rosenbrock, D=20, HMC)",
                         "peak_source": "mcmc_fp64_peak_probe measured in this run (register-resident DFMA, all SMs); "
                                        "MEASURED_PEAKS.json has no fp64 entry; nominal 37.2 TFLOP/s",
                         "flop_model": f"{F_STEP:.0f} flops per leapfrog step + {F_ITER:.0f} per MH iteration (FMA=2), "
                                       "this kernel's own formulation; RNG excluded",
                         "writeback": {"bytes_per_call": wb_bytes, "unit": "GB/s",
                                       "achieved_if_alone": {str(L): wb_bytes / (kern_ms[L] / args.steps * 1e-3) / 1e9
                                                             for L in L_GRID}}},
            "e2e": {"value": e2e_value, "unit": "grad-evals/s", "h2d_bytes_per_step": h2d * len(L_GRID),
                    "d2h_bytes_per_step": d2h * len(L_GRID), "steps": e2e_steps,
                    "what": "hmc_run(pinned host init) -> samples, log_probs, accept_rate, final state copied to pinned host"},
            "gpu_launches": len(L_GRID) * args.steps,
            "min_ess_per_sec": {str(L): ess[L]["min_ess_per_sec"] for L in L_GRID},
            "ess": {str(L): ess[L] for L in L_GRID},
            "comm": comm,
            "clocks": clocks,
            "mean_accept_rate_last_call": acc_sum,
        }
        if world == 1 and not args.no_cpu_baseline:
            cpu_pass(2048, 8)  # wake the OpenMP team
            r = cpu_pass(C, S_STORED)   # one whole step of the workload: about 4-5 s on 16 threads
            line["cpu_baseline"] = {"value": r["value"], "unit": "grad-evals/s", "cores": r["cores"], "kind": "port",
                                    "sample": f"one whole step: all {C} chains x {S_STORED} stored iterations x L grid "
                                              f"{list(L_GRID)} ({r['seconds']:.1f} s)",
                                    "note": "reference JAX CPU path unavailable on this box (jax not installed); C "
                                            "restatement oracle/mcmc_oracle.c, OpenMP over chains, gcc " + r["flags"]}
        emit_json(line)
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    ensure_built()
    if args.impl == "reference":
        run_reference_arm(args)
    else:
        run_gpu_arm(args)


if __name__ == "__main__":
    main()
