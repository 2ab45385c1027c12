"""Stage timings of mcmc_b200.diagnostics.compute_diagnostics at the config-5 shape (65,536 chains x 256 draws x 20 dims),
so that the 0.75 s of round 1 can be attributed: python tools/diag_probe.py [--chains C] [--draws S]."""
import argparse
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from mcmc_b200 import diagnostics as DG  # noqa: E402


def t(fn, reps=3):
    fn()
    torch.cuda.synchronize()
    best = 1e9
    for _ in range(reps):
        t0 = time.perf_counter()
        r = fn()
        torch.cuda.synchronize()
        best = min(best, time.perf_counter() - t0)
    return best * 1e3, r


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--chains", type=int, default=65536)
    ap.add_argument("--draws", type=int, default=256)
    a = ap.parse_args()
    D = 20
    g = torch.Generator(device="cuda").manual_seed(0)
    x = torch.randn((a.draws, a.chains, D), dtype=torch.float64, device="cuda", generator=g)
    x[1:] = 0.7 * x[:-1] + 0.3 * x[1:]
    ms, _ = t(lambda: DG.compute_diagnostics(x), reps=2)
    print(f"compute_diagnostics total: {ms:.1f} ms")
    col = x[:, :, 3]
    ms, (split, n) = t(lambda: DG._split_rows(col.to(torch.float64)))
    print(f"  split_rows (strided column -> contiguous): {ms:.2f} ms")
    flat = split.reshape(-1).contiguous()
    ms, (vals, perm) = t(lambda: torch.sort(flat))
    print(f"  torch.sort {flat.numel()/1e6:.1f} M doubles: {ms:.2f} ms")
    ms, (z, _) = t(lambda: DG.rank_normalize(split))
    print(f"  rank_normalize (sort + kernel): {ms:.2f} ms")
    ms, _ = t(lambda: DG.rank_normalize_fold(split))
    print(f"  rank_normalize_fold (one pair sort with u32 indices + fused bulk/fold rank kernel): {ms:.2f} ms")
    ms, _ = t(lambda: DG.split_chain_stats(z, n, n, None))
    print(f"  split_chain_stats all lags: {ms:.2f} ms")
    ms, _ = t(lambda: DG.split_chain_stats(z, n, n, 1))
    print(f"  split_chain_stats 1 lag: {ms:.2f} ms")
    ms, _ = t(lambda: (split <= 0.3).to(torch.float64))
    print(f"  indicator: {ms:.2f} ms")
    ms, _ = t(lambda: (flat.mean(), flat.std()))
    print(f"  mean/std: {ms:.2f} ms")
    ms, _ = t(lambda: DG.variable_diagnostics(col))
    print(f"  variable_diagnostics (one variable): {ms:.2f} ms")
    ms, _ = t(lambda: torch.sort(x.reshape(-1, D).T.contiguous(), dim=1))
    print(f"  batched sort of all {D} variables: {ms:.2f} ms")
