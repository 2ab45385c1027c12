"""world_size-2 gloo tests of the multi-GPU host logic (mcmc_b200/dist.py and the sharded warmup):
chain sharding + the cross-chain reductions reproduce the single-process result."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as td
import torch.multiprocessing as mp


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _spawn(fn, world, *args):
    port = _free_port()
    ctx = mp.get_context("spawn")
    q = ctx.SimpleQueue()
    procs = [ctx.Process(target=_entry, args=(fn, r, world, port, q) + args) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(300)
    res = {}
    while not q.empty():
        r, v = q.get()
        res[r] = v
    for p in procs:
        assert p.exitcode == 0, f"worker exit code {p.exitcode}"
    return res


def _entry(fn, rank, world, port, q, *args):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world),
                      OMP_NUM_THREADS="2")
    td.init_process_group("gloo", rank=rank, world_size=world)
    try:
        q.put((rank, fn(rank, world, *args)))
    finally:
        td.destroy_process_group()


def _reductions(rank, world):
    from mcmc_b200 import dist
    out = {}
    dist.warm_up()                                   # every collective once (gloo: the all-to-all is skipped)
    out["shard"] = dist.chain_shard(11)
    a, b = dist.fused_sum([torch.tensor([1.0 + rank, 2.0]), torch.tensor([[3 + rank]], dtype=torch.int64)])
    out["fused"] = (a.tolist(), b.tolist())
    out["mean_accept"] = dist.mean_accept(torch.tensor(10.0 * (rank + 1)), 100 * (rank + 1))
    out["var"] = dist.mean_chain_variance(torch.tensor([4.0, 8.0]) * (rank + 1), 2 * (rank + 1)).tolist()
    x = torch.arange((3 + rank) * 2, dtype=torch.float64).reshape(3 + rank, 2) + 100 * rank
    out["gather"] = dist.gather_chains(x).tolist()
    out["gather_cap"] = dist.gather_chains(x, max_per_rank=2).shape[0]
    t = torch.tensor([float(rank)])
    out["max"] = dist.all_reduce_max_(t).item()
    return out


def test_reductions_world2():
    r = _spawn(_reductions, 2)
    assert r[0]["shard"] == (0, 6) and r[1]["shard"] == (6, 5)
    for k in (0, 1):
        assert r[k]["fused"] == ([3.0, 4.0], [[7.0]])
        assert r[k]["mean_accept"] == pytest.approx(30.0 / 300.0)
        assert r[k]["var"] == pytest.approx([12.0 / 6, 24.0 / 6])
        assert len(r[k]["gather"]) == 7 and r[k]["gather"][3] == [100.0, 101.0]
        assert r[k]["gather_cap"] == 4 and r[k]["max"] == 1.0


def test_single_process_is_identity():
    from mcmc_b200 import dist
    assert dist.world_size() == 1 and dist.rank() == 0 and dist.chain_shard(9) == (0, 9)
    t = torch.tensor([1.0, 2.0])
    assert dist.fused_sum([t])[0].tolist() == [1.0, 2.0]
    assert dist.mean_accept(torch.tensor(5.0), 10) == 0.5
    assert dist.gather_chains(t.reshape(2, 1)).shape == (2, 1)


def _warmup(rank, world, total_chains, D):
    from mcmc_b200 import dist
    from mcmc_b200.benchmarks import targets as T
    from mcmc_b200.tuning import run_adaptive_warmup
    from helpers import patch_tuning_to_oracle
    patch_tuning_to_oracle()
    tgt = T.neals_funnel(D)
    q_all = np.random.default_rng(5).standard_normal((total_chains, D))
    off, n = dist.chain_shard(total_chains)
    eps, w, pos, info = run_adaptive_warmup("grahmc", tgt.log_prob_fn, None, q_all[off:off + n], key=11,
                                            num_warmup=2500, schedule_type="sine", num_steps=6, chain_offset=off,
                                            tune_friction=False, verbose=False)
    return dict(eps=eps, w=w.numpy(), pos=pos.numpy(), off=off)


def test_sharded_warmup_equals_unsharded():
    """GRAHMC warmup on the funnel, 24 chains: 2 gloo ranks x 12 chains reproduce the 1-rank run —
    same Philox streams (chain_offset), accept sums exact, variance sums equal up to fp64 reduction order."""
    C, D = 24, 5
    two = _spawn(_warmup, 2, C, D)
    one = _spawn(_warmup, 1, C, D)[0]
    assert two[0]["eps"] == two[1]["eps"]
    np.testing.assert_allclose(two[0]["w"], two[1]["w"], rtol=0, atol=0)
    assert two[0]["eps"] == pytest.approx(one["eps"], rel=1e-9)
    # a 1-ulp difference in the reduced variances feeds the next window's trajectories (funnel: mildly chaotic)
    np.testing.assert_allclose(two[0]["w"], one["w"], rtol=1e-6)
    pos2 = np.concatenate([two[0]["pos"], two[1]["pos"]])
    print("max |pos diff|", np.abs(pos2 - one["pos"]).max())
    np.testing.assert_allclose(pos2, one["pos"], rtol=1e-3, atol=1e-3)
