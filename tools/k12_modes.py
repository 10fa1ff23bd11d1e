#!/usr/bin/env python
"""One strand of the bench workload (cfg2, 402 M positions) on one GPU, kernel by kernel: the fused kernel with direct
adds (lime_strand_run) against the fused kernel appending records + routing + applying them (lime_xchg_*, loopback,
world = 1 and world = 8 virtual owners).  Prints ms per launch of every kernel family (CUDA events around each launch).
    python tools/k12_modes.py [n_positions] [alpha]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import torch

from lime_b200 import datagen
from lime_b200.api import LimeContext, MEM_DEVICE

n = int(sys.argv[1]) if len(sys.argv) > 1 else 402_000_000
alpha = int(sys.argv[2]) if len(sys.argv) > 2 else 16
n_reads, n_refs, read_len = max(1000, int(2_000_000 * n / 402_000_000)), 100, 100
dev = torch.device("cuda", 0)
e, l, d = datagen.synth_arrays(n, n_reads, n_refs, read_len, seed=20261106, device=dev)
torch.cuda.synchronize()
v = np.arange(n_refs, dtype=np.uint32)


def show(tag, prof, reps):
    print(tag, {k: round(x["ms"] / x["launches"], 3) for k, x in prof.items() if x["launches"]},
          "launches/rep", {k: x["launches"] // reps for k, x in prof.items() if x["launches"]}, flush=True)


ctx = LimeContext(0)
ctx.set_stream(torch.cuda.current_stream().cuda_stream)
reps = 5
for rep in range(reps + 2):
    if rep == 2:
        ctx.prof_enable(True); ctx.prof_reset()
    for s in range(2):
        ctx.load(l.data_ptr(), d.data_ptr(), e.data_ptr(), n_reads, n_refs, mem=MEM_DEVICE, n_total=n)
        ctx.strand_run(s, alpha, 1, read_len + 1 - alpha, 0.25)
    ctx.classify(2, n_refs, v, None, 1, False)
show("direct:", ctx.prof(), reps)
st = ctx.strand_status(0)
print("strand:", dict(zip(("n_clusters", "max_len", "n_records", "updates"), st)), flush=True)
ctx.prof_enable(False)
ctx.close()

for world in (1, 8):
    ctx = LimeContext(0)
    ctx.set_stream(torch.cuda.current_stream().cuda_stream)
    per = (n_reads + world - 1) // world
    flat_cap = int(n * 0.25) + (1 << 21)
    ctx.xchg_alloc(world, 0, per, n_refs, 2, flat_cap, int(flat_cap * 1.5 / world) + (1 << 16))
    ctx.xchg_open(None)
    for rep in range(reps + 2):
        if rep == 2:
            ctx.prof_enable(True); ctx.prof_reset()
        for s in range(2):
            ctx.load(l.data_ptr(), d.data_ptr(), e.data_ptr(), n_reads, n_refs, mem=MEM_DEVICE, n_total=n)
            ctx.xchg_accumulate(alpha, 1, s)
            ctx.xchg_apply(s)
            ctx.xchg_choose(s, read_len + 1 - alpha, 0.25, min(per, n_reads))
        ctx.classify(2, n_refs, v, None, 1, False)
    show(f"records, {world} owner(s):", ctx.prof(), reps)
    print("  status:", ctx.xchg_status(0), flush=True)
    ctx.prof_enable(False)
    ctx.close()
