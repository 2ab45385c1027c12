"""Two-GPU NCCL checks of the chain-sharded path (skipped on a single-GPU box): sharded sampling reproduces the
single-GPU run bit for bit, and the distributed diagnostics / warmup equal the single-process results."""
import os
import socket
import subprocess
import sys

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

WORKER = r'''
import os, sys, json
import numpy as np, torch, torch.distributed as td
sys.path.insert(0, os.environ["MCMC_ROOT"])
from mcmc_b200 import dist, diagnostics as DG
from mcmc_b200.benchmarks import targets as T
from mcmc_b200.benchmarks.metrics import compute_sliced_w2
from mcmc_b200.random import normal
from mcmc_b200.samplers import hmc_run
from mcmc_b200.tuning import run_adaptive_warmup
local = int(os.environ["LOCAL_RANK"]); torch.cuda.set_device(local)
td.init_process_group("nccl", device_id=torch.device("cuda", local))
C, D = 4096, 20
tgt = T.correlated_gaussian(D, 0.5)
q_all = normal(3, (C, D))
off, n = dist.chain_shard(C)
out = hmc_run(9, tgt.log_prob_fn, torch.from_numpy(q_all[off:off + n]).cuda(), 0.3, 6, 64, burn_in=16, _chain_offset=off)
d = DG.compute_diagnostics(out[0], distributed=True)
sw2 = compute_sliced_w2(out[0], "correlated_gaussian", D, n_reference=20000, distributed=True, correlation=0.5)
ill = T.ill_conditioned_gaussian(D, 25.0)
eps, w, pos, info = run_adaptive_warmup("hmc", ill.log_prob_fn, None, q_all[off:off + n], key=4, num_warmup=2500,
                                        num_steps=8, chain_offset=off, verbose=False)
np.save(os.environ["MCMC_OUT"] + f"/samples_{dist.rank()}.npy", out[0].cpu().numpy())
if dist.rank() == 0:
    json.dump({"rhat": d["rhat"].tolist(), "ess_bulk": d["ess_bulk"].tolist(), "ess_tail": d["ess_tail"].tolist(),
               "mcse": d["summary"]["mcse_mean"].tolist(), "sw2": sw2, "eps": eps, "w": w.cpu().tolist()},
              open(os.environ["MCMC_OUT"] + "/dist.json", "w"))
td.destroy_process_group()
'''


def _port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def test_two_gpu_shards_match_single_gpu(gpu_lib, tmp_path):
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs (gpurun --gpus 2)")
    import json
    from mcmc_b200 import diagnostics as DG
    from mcmc_b200.benchmarks import targets as T
    from mcmc_b200.random import normal
    from mcmc_b200.samplers import hmc_run
    from mcmc_b200.tuning import run_adaptive_warmup
    script = tmp_path / "worker.py"
    script.write_text(WORKER)
    env = dict(os.environ, MCMC_ROOT=ROOT, MCMC_OUT=str(tmp_path))
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2",
                        "--master-addr", "127.0.0.1", "--master-port", str(_port()), str(script)],
                       env=env, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-3000:]
    C, D = 4096, 20
    tgt = T.correlated_gaussian(D, 0.5)
    q_all = normal(3, (C, D))
    out = hmc_run(9, tgt.log_prob_fn, torch.from_numpy(q_all).cuda(), 0.3, 6, 64, burn_in=16)
    both = np.concatenate([np.load(tmp_path / "samples_0.npy"), np.load(tmp_path / "samples_1.npy")], axis=1)
    assert np.array_equal(both, out[0].cpu().numpy())           # Philox keyed on the global chain id
    got = json.load(open(tmp_path / "dist.json"))
    d = DG.compute_diagnostics(out[0])
    np.testing.assert_allclose(got["rhat"], d["rhat"], rtol=1e-12)
    np.testing.assert_allclose(got["ess_bulk"], d["ess_bulk"], rtol=1e-10)
    np.testing.assert_allclose(got["ess_tail"], d["ess_tail"], rtol=1e-10)
    np.testing.assert_allclose(got["mcse"], d["summary"]["mcse_mean"], rtol=1e-10)
    assert 0.0 < got["sw2"] < 0.1
    ill = T.ill_conditioned_gaussian(D, 25.0)
    eps, w, pos, info = run_adaptive_warmup("hmc", ill.log_prob_fn, None, q_all, key=4, num_warmup=2500, num_steps=8,
                                            verbose=False)
    assert got["eps"] == pytest.approx(eps, rel=1e-6)
    np.testing.assert_allclose(got["w"], w.cpu().numpy(), rtol=1e-6)
