"""Host <-> device copy ceiling of the box with ALL ranks copying at once: what bench.py's `e2e` number (ids +
gradients up, embeddings down, every step, every rank) can reach at best on this host.  One rank per GPU:

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 tools/pcie_probe.py
    python tools/pcie_probe.py                       # one GPU

Each rank moves one step's worth of the cfg2 e2e traffic (436 MB up and 436 MB down) between pinned host
memory and its GPU: up alone, down alone, both directions at once; timed between barriers, max over ranks.
Prints one JSON line: per-rank and aggregate GB/s, and where the host's CPUs / NUMA node sit for each GPU.
"""
import json
import os
import time

import torch
import torch.distributed as dist

rank = int(os.environ.get("RANK", "0"))
world = int(os.environ.get("WORLD_SIZE", "1"))
local = int(os.environ.get("LOCAL_RANK", "0"))
dev = torch.device(f"cuda:{local}")
torch.cuda.set_device(dev)
if world > 1:
    dist.init_process_group("nccl", device_id=dev)
n = 436 * 1024 * 1024 // 4
h_up = torch.empty(n, dtype=torch.float32).pin_memory()
h_dn = torch.empty(n, dtype=torch.float32).pin_memory()
d_up = torch.empty(n, dtype=torch.float32, device=dev)
d_dn = torch.empty(n, dtype=torch.float32, device=dev)
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()


def barrier():
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()


def timed(f, k=5):
    f()
    barrier()
    t0 = time.perf_counter()
    for _ in range(k):
        f()
    torch.cuda.synchronize()
    dt = (time.perf_counter() - t0) / k
    t = torch.tensor([dt], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def up():
    d_up.copy_(h_up, non_blocking=True)


def down():
    h_dn.copy_(d_dn, non_blocking=True)


def both():
    with torch.cuda.stream(s1):
        d_up.copy_(h_up, non_blocking=True)
    with torch.cuda.stream(s2):
        h_dn.copy_(d_dn, non_blocking=True)


gb = n * 4 / 1e9
res = {}
for name, f, factor in (("h2d", up, 1), ("d2h", down, 1), ("both_directions", both, 2)):
    s = timed(f)
    res[name] = {"ms": round(s * 1e3, 2), "per_rank_gbs": round(factor * gb / s, 1),
                 "aggregate_gbs": round(world * factor * gb / s, 1)}
numa = None
try:
    bus = torch.cuda.get_device_properties(dev).pci_bus_id
    domain = torch.cuda.get_device_properties(dev).pci_domain_id
    path = f"/sys/bus/pci/devices/{domain:04x}:{bus:02x}:00.0/numa_node"
    numa = int(open(path).read())
except Exception:
    pass
info = [None] * world
mine = {"rank": rank, "numa_node": numa, "cpus_allowed": len(os.sched_getaffinity(0))}
if world > 1:
    dist.all_gather_object(info, mine)
else:
    info = [mine]
if rank == 0:
    print(json.dumps({"probe": "pinned host <-> device copies, all ranks at once", "ranks": world,
                      "bytes_per_copy": n * 4, "results": res, "ranks_info": info}), flush=True)
if world > 1:
    dist.destroy_process_group()
