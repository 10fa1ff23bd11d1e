#!/usr/bin/env python
"""Probe of the inter-GPU path of a multi-GPU box: topology, peer access, device-to-device copy
bandwidth, and (under torchrun) NCCL all-to-all bandwidth with the transports NCCL chose."""
import os, subprocess, sys, time
import torch

if "RANK" not in os.environ:
    print(subprocess.run(["nvidia-smi", "topo", "-m"], capture_output=True, text=True).stdout)
    n = torch.cuda.device_count()
    for a in range(n):
        for b in range(n):
            if a != b:
                print(f"can_device_access_peer({a},{b}) = {torch.cuda.can_device_access_peer(a, b)}")
    if n >= 2:
        x = torch.empty(1 << 28, dtype=torch.uint8, device="cuda:0"); y = torch.empty_like(x, device="cuda:1")
        for _ in range(2): y.copy_(x)
        torch.cuda.synchronize(0); torch.cuda.synchronize(1)
        t = time.time()
        for _ in range(5): y.copy_(x)
        torch.cuda.synchronize(0); torch.cuda.synchronize(1)
        print(f"cuda:0 -> cuda:1 copy: {5 * x.numel() / (time.time() - t) / 1e9:.1f} GB/s")
    sys.exit(0)

import torch.distributed as dist
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
n = 1 << 24                                   # int64 elements per peer (128 MB)
send = torch.ones(n * world, dtype=torch.int64, device="cuda"); recv = torch.empty_like(send)
for _ in range(3): dist.all_to_all_single(recv, send)
torch.cuda.synchronize(); dist.barrier()
t = time.time()
for _ in range(5): dist.all_to_all_single(recv, send)
torch.cuda.synchronize()
dt = (time.time() - t) / 5
if rank == 0:
    print(f"NCCL all_to_all: {n * 8 * (world - 1) / dt / 1e9:.1f} GB/s sent per rank ({n * 8 / 1e6:.0f} MB per peer, {dt * 1e3:.2f} ms)")
dist.destroy_process_group()
