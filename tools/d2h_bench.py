"""Plain pinned device->host copy bandwidth at N ranks (one process per GPU), to tell the box's ceiling from the library's
copy path:  python tools/d2h_bench.py   or   torchrun --nproc-per-node N tools/d2h_bench.py
Every rank copies one 1.5 GB device buffer into one pinned host buffer, 5 times; rank 0 prints per-rank and total GB/s."""
import json
import os
import time

import torch
import torch.distributed as dist

rank = int(os.environ.get("RANK", "0"))
world = int(os.environ.get("WORLD_SIZE", "1"))
local = int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from mazetd_b200 import numa

numa_info = numa.bind_to_gpu_node(local) if world > 1 else None  # MTD_NO_NUMA_BIND=1 leaves the affinity alone
if world > 1:
    dist.init_process_group("nccl", device_id=dev)
n = 1536 * 1024 * 1024
d = torch.empty(n, dtype=torch.uint8, device=dev)
h = torch.empty(n, dtype=torch.uint8, pin_memory=True)
h.copy_(d, non_blocking=True)
torch.cuda.synchronize()
res = []
for it in range(5):
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    h.copy_(d, non_blocking=True)
    torch.cuda.synchronize()
    res.append(n / (time.perf_counter() - t0) / 1e9)
best = max(res)
t = torch.tensor([best], dtype=torch.float64, device=dev)
parts = [torch.zeros_like(t) for _ in range(world)]
if world > 1:
    dist.all_gather(parts, t)
else:
    parts = [t]
if rank == 0:
    per = [float(p.item()) for p in parts]
    print(json.dumps({"n_gpus": world, "bytes_per_rank": n, "d2h_GBps_per_rank": [round(x, 1) for x in per], "d2h_GBps_total": round(sum(per), 1), "host_cpus": len(os.sched_getaffinity(0)), "numa_rank0": numa_info}), flush=True)
if world > 1:
    dist.destroy_process_group()
