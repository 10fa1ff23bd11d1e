#!/usr/bin/env python
"""Rate of random 32-bit adds into peer-mapped similarity slabs (lime_peer_probe) under torchrun:
local vs remote, footprint (rows touched), red / atom / store, one rank active vs all ranks."""
import os, sys
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from lime_b200.api import LimeContext

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
ctx = LimeContext(local)
ctx.set_stream(torch.cuda.current_stream().cuda_stream)
ROWS = int(os.environ.get("PROBE_ROWS", (16 << 20) if world == 2 else (2 << 20)))
h = ctx.peer_alloc(ROWS, 100, 1)
mine = torch.from_numpy(h.copy()).to(dev)
parts = [torch.empty_like(mine) for _ in range(world)]
dist.all_gather(parts, mine)
ctx.peer_open(world, rank, torch.stack(parts).cpu().numpy(), ROWS)
dist.barrier()
N = 40_000_000
nxt = (rank + 1) % world


def run(tag, first, count, span, mode, everyone):
    dist.barrier(); torch.cuda.synchronize()
    ms = 0.0
    if everyone or rank == 0:
        ctx.peer_probe(0, first, count, span, N // 8, mode)
        ms = ctx.peer_probe(0, first, count, span, N, mode)
    t = torch.tensor([ms], device=dev, dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    if rank == 0:
        print(f"{tag:58s} {float(t.item()):9.3f} ms  {N / float(t.item()) / 1e6:8.2f} G adds/s per GPU", flush=True)


for everyone in (False, True):
    who = "all ranks" if everyone else "rank 0 only"
    run(f"[{who}] local red, 2M rows (224 MB)", rank, 1, 2 << 20, 0, everyone)
    for span in [x for x in (1 << 18, 2 << 20, 8 << 20, 16 << 20) if x <= ROWS]:
        run(f"[{who}] remote red -> next rank, {span >> 10}K rows ({span * 112 >> 20} MB)", nxt, 1, span, 0, everyone)
    run(f"[{who}] remote atom -> next rank, 2M rows", nxt, 1, 2 << 20, 1, everyone)
    run(f"[{who}] remote store -> next rank, 2M rows", nxt, 1, 2 << 20, 2, everyone)
    if world > 2:
        if not everyone:
            for k in range(1, world):
                run(f"[{who}] red -> rank {k}, 2M rows", k, 1, 2 << 20, 0, everyone)
        for cnt in (2, 4, world):
            if cnt <= world:
                run(f"[{who}] red -> owners 0..{cnt - 1}, 2M rows each", 0, cnt, 2 << 20, 0, everyone)
        run(f"[{who}] red -> all {world} owners, 256K rows each", 0, world, 1 << 18, 0, everyone)
        run(f"[{who}] store -> all {world} owners, 2M rows each", 0, world, 2 << 20, 2, everyone)
        run(f"[{who}] red.add 1 -> all {world} owners, 2M rows each", 0, world, 2 << 20, 3, everyone)
    run(f"[{who}] red.add 1 -> next rank, 2M rows", nxt, 1, 2 << 20, 3, everyone)
dist.barrier()
ctx.peer_close()
ctx.close()
dist.destroy_process_group()
