"""Concurrent device-to-host bandwidth of all ranks of one box (what bounds the
end-to-end number at N > 1): each rank copies 2 GiB from its GPU into pinned host
memory, all ranks at once, (a) as launched and (b) with the process bound to the CPUs
NVML reports as local to its GPU before the pinned buffer is allocated.  Prints the
topology facts the numbers depend on.  Run on the GPU box:
    python -m torch.distributed.run --nproc-per-node 8 --master-addr 127.0.0.1 \
        profiles/tools/d2h_scaling.py"""
import json
import os
import subprocess
import time

import torch
import torch.distributed as dist

rank = int(os.environ.get("RANK", "0"))
local = int(os.environ.get("LOCAL_RANK", "0"))
world = int(os.environ.get("WORLD_SIZE", "1"))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
if world > 1:
    dist.init_process_group("nccl", device_id=dev)


def gpu_local_cpus(index):
    import pynvml
    pynvml.nvmlInit()
    h = pynvml.nvmlDeviceGetHandleByIndex(index)
    n = os.cpu_count()
    words = pynvml.nvmlDeviceGetCpuAffinity(h, (n + 63) // 64)
    return [c for c in range(n) if (words[c // 64] >> (c % 64)) & 1]


def measure(tag):
    n = 2 << 30
    src = torch.empty(n, dtype=torch.uint8, device=dev)
    dst = torch.empty(n, dtype=torch.uint8).pin_memory()
    for _ in range(2):
        dst.copy_(src, non_blocking=True)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    t0 = time.perf_counter()
    reps = 5
    for _ in range(reps):
        dst.copy_(src, non_blocking=True)
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    gbs = reps * n / dt / 1e9
    t = torch.tensor([gbs], dtype=torch.float64, device=dev)
    if world > 1:
        g = [torch.zeros_like(t) for _ in range(world)]
        dist.all_gather(g, t)
        vals = [float(x.item()) for x in g]
    else:
        vals = [gbs]
    if rank == 0:
        print(json.dumps({tag: {"per_rank_gbs": [round(v, 1) for v in vals],
                                "sum_gbs": round(sum(vals), 1)}}))
    del src, dst


if rank == 0:
    print(json.dumps({"cpus": os.cpu_count(), "world": world}))
    try:
        print(subprocess.run(["nvidia-smi", "topo", "-m"], capture_output=True, text=True,
                             timeout=30).stdout[-3000:])
    except Exception as e:  # noqa: BLE001
        print("topo failed", e)
measure("as_launched")
try:
    cpus = gpu_local_cpus(local)
    if cpus:
        os.sched_setaffinity(0, cpus)
    if rank == 0:
        print(json.dumps({"rank0_local_cpus": cpus[:8] + (["..."] if len(cpus) > 8 else []),
                          "count": len(cpus)}))
except Exception as e:  # noqa: BLE001
    if rank == 0:
        print("affinity failed", e)
measure("bound_to_gpu_local_cpus")
if world > 1:
    dist.destroy_process_group()
