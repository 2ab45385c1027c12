"""Three calls of diagnostics.rank_normalize_fold at the config-5 size (256 draws x 65,536 chains of one variable), for an
ncu launch list of its kernels:
    ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/fold_launches.csv python tools/fold_probe.py"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from mcmc_b200 import diagnostics as DG  # noqa: E402

g = torch.Generator(device="cuda").manual_seed(0)
x = torch.randn((256, 65536), dtype=torch.float64, device="cuda", generator=g)
for _ in range(3):
    DG.rank_normalize_fold(x)
torch.cuda.synchronize()
