"""Device-to-host bandwidth of the box, one rank per GPU (torchrun): what bounds bench.py's `e2e` at N GPUs.

For every rank: the GPU's PCI address and NUMA node (sysfs), the CPUs the process may run on, then D2H copies of a
1 GiB buffer into pinned host memory, timed with CUDA events
  solo        ranks take turns, one GPU copying at a time
  concurrent  all ranks copy at once (what the e2e leg of bench.py does at N GPUs)
  bound       the same after binding the process to the CPUs of the GPU's NUMA node and re-allocating the pinned buffer
              (first-touch placement next to the GPU), if sysfs exposes the node
Rank 0 prints one JSON line per test with the per-rank GB/s and their sum.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511 tools/d2h_probe.py
"""
import json
import os
import sys

import torch
import torch.distributed as dist

GIB = 1 << 30
REPS = 6


def sysfs(path):
    try:
        return open(path).read().strip()
    except OSError:
        return None


def cpulist(s):
    out = []
    for part in (s or "").split(","):
        if "-" in part:
            a, b = part.split("-")
            out += list(range(int(a), int(b) + 1))
        elif part:
            out.append(int(part))
    return out


def timed_copies(dst, src, barrier_first):
    if barrier_first:
        dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(REPS):
        dst.copy_(src, non_blocking=True)
    e1.record()
    torch.cuda.synchronize()
    return REPS * GIB / (e0.elapsed_time(e1) * 1e-3) / 1e9


def gather(x):
    t = torch.tensor([x], dtype=torch.float64, device="cuda")
    out = [torch.zeros_like(t) for _ in range(dist.get_world_size())]
    dist.all_gather(out, t)
    return [float(o) for o in out]


def main():
    rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    props = torch.cuda.get_device_properties(local)
    bdf = None
    try:
        bdf = f"{props.pci_domain_id:04x}:{props.pci_bus_id:02x}:{props.pci_device_id:02x}.0"
    except AttributeError:
        pass
    node = sysfs(f"/sys/bus/pci/devices/{bdf}/numa_node") if bdf else None
    info = dict(rank=rank, gpu=local, pci=bdf, numa_node=node, affinity=len(os.sched_getaffinity(0)),
                nodes_online=sysfs("/sys/devices/system/node/online"))
    infos = [None] * world
    dist.all_gather_object(infos, info)
    if rank == 0:
        print(json.dumps(dict(test="topology", ranks=infos)), flush=True)

    src = torch.empty(GIB, dtype=torch.uint8, device="cuda").fill_(1)
    dst = torch.empty(GIB, dtype=torch.uint8).pin_memory()
    dst.copy_(src)   # touch
    torch.cuda.synchronize()

    solo = 0.0
    for r in range(world):
        dist.barrier()
        if r == rank:
            solo = timed_copies(dst, src, False)
    dist.barrier()
    res = gather(solo)
    if rank == 0:
        print(json.dumps(dict(test="solo", gbs=[round(x, 1) for x in res])), flush=True)
    conc = gather(timed_copies(dst, src, True))
    if rank == 0:
        print(json.dumps(dict(test="concurrent", gbs=[round(x, 1) for x in conc], total=round(sum(conc), 1))), flush=True)

    cpus = cpulist(sysfs(f"/sys/devices/system/node/node{node}/cpulist")) if node not in (None, "-1") else []
    cpus = sorted(set(cpus) & os.sched_getaffinity(0))
    bound_ok = bool(cpus)
    if bound_ok:
        os.sched_setaffinity(0, cpus)
        del dst
        dst = torch.empty(GIB, dtype=torch.uint8).pin_memory()
        dst.copy_(src)
        torch.cuda.synchronize()
    b = gather(timed_copies(dst, src, True))
    oks = gather(1.0 if bound_ok else 0.0)
    if rank == 0:
        print(json.dumps(dict(test="bound", gbs=[round(x, 1) for x in b], total=round(sum(b), 1), ranks_bound=int(sum(oks)))), flush=True)
    # host to device for completeness (the e2e leg uploads 42 MB per step: negligible next to the 2.9 GB of results)
    torch.cuda.synchronize()
    h2d = gather(timed_copies(src, dst, True))
    if rank == 0:
        print(json.dumps(dict(test="h2d_concurrent", gbs=[round(x, 1) for x in h2d], total=round(sum(h2d), 1))), flush=True)
    dist.destroy_process_group()


if __name__ == "__main__":
    sys.exit(main())
