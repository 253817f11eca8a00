"""D2H copy rate of one rank alone and of all ranks at once (torchrun): the e2e leg moves 4.46 MB per step and rank as
four 1.1 MB copies.  Prints GB/s per rank for both cases, and the host cost of an empty graph replay + event sync."""
import os
import time

import torch
import torch.distributed as dist

rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
local = int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
if world > 1:
    dist.init_process_group("nccl", device_id=dev)
nbytes = 1114112  # one shard's block
src = torch.empty((nbytes,), dtype=torch.uint8, device=dev)
dst = [torch.empty((nbytes,), dtype=torch.uint8).pin_memory() for _ in range(4)]


def rate(n=400):
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for i in range(n):
        dst[i & 3].copy_(src, non_blocking=True)
    torch.cuda.synchronize()
    return n * nbytes / (time.perf_counter() - t0) / 1e9


def sync():
    if world > 1:
        dist.barrier()


rate(50)
alone = None
for r in range(world):  # one rank at a time
    sync()
    if r == rank:
        alone = rate()
sync()
together = rate()
g = torch.cuda.CUDAGraph()
x = torch.zeros(8, device=dev)
with torch.cuda.graph(g):
    x.add_(1)
ev = torch.cuda.Event()
torch.cuda.synchronize()
t0 = time.perf_counter()
for _ in range(2000):
    g.replay()
    ev.record()
    ev.synchronize()
host_us = (time.perf_counter() - t0) / 2000 * 1e6
out = torch.tensor([alone, together, host_us], dtype=torch.float64, device=dev)
if world > 1:
    allv = [torch.zeros_like(out) for _ in range(world)]
    dist.all_gather(allv, out)
else:
    allv = [out]
if rank == 0:
    for r, v in enumerate(allv):
        print(f"rank {r}: D2H alone {v[0].item():.1f} GB/s, all ranks at once {v[1].item():.1f} GB/s, graph replay + event sync {v[2].item():.1f} us")
    print("affinity", sorted(os.sched_getaffinity(0)), "cpus", os.cpu_count())
if world > 1:
    dist.destroy_process_group()
