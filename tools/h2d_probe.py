#!/usr/bin/env python
"""Platform ceiling for concurrent pinned H2D copies: every rank copies the same amount the e2e
bench moves per step.  torchrun --nproc-per-node N tools/h2d_probe.py"""
import os
import time

import torch
import torch.distributed as dist

rank, world, lr = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(lr)
if world > 1:
    dist.init_process_group("nccl", device_id=torch.device("cuda", lr))
n = 688_128_000
h = torch.empty(n, dtype=torch.uint8).pin_memory()
d = torch.empty(n, dtype=torch.uint8, device="cuda")
for _ in range(2):
    d.copy_(h, non_blocking=True)
torch.cuda.synchronize()
if world > 1:
    dist.barrier()
t0 = time.perf_counter()
for _ in range(10):
    d.copy_(h, non_blocking=True)
torch.cuda.synchronize()
dt = time.perf_counter() - t0
gbs = torch.tensor([n * 10 / dt / 1e9], device="cuda")
if world > 1:
    out = [torch.zeros_like(gbs) for _ in range(world)]
    dist.all_gather(out, gbs)
    if rank == 0:
        v = [float(x) for x in out]
        print("H2D GB/s per rank:", ["%.1f" % x for x in v], "sum %.1f" % sum(v))
    dist.destroy_process_group()
else:
    print("H2D GB/s: %.1f" % float(gbs))
