#!/usr/bin/env python
"""Diagnostic (torchrun, >= 2 GPUs): the record exchange step by step against the single-GPU matrix.
Each rank holds the whole (small) arrays, scores ITS position shard with lime_xchg_accumulate, the counts are
all-gathered, the owner applies; the slab must equal the owner's rows of the single-GPU matrix."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import torch
import torch.distributed as dist

from lime_b200 import datagen, multi
from lime_b200.api import LimeContext, MEM_DEVICE

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
n, n_reads, n_refs, alpha, halo = int(os.environ.get('DBG_N', 2_000_003)), int(os.environ.get('DBG_READS', 20_000)), 40, 16, int(os.environ.get('DBG_HALO', 70_000))
e, l, d = datagen.synth_arrays(n, n_reads, n_refs, 100, seed=77, device=dev, p_long=2e-5, max_seg=60000)
torch.cuda.synchronize()
ref = LimeContext(local)
ref.set_stream(torch.cuda.current_stream().cuda_stream)
ref.load(l.data_ptr(), d.data_ptr(), e.data_ptr(), n_reads, n_refs, mem=MEM_DEVICE, n_total=n)
ref.cluster_similarity(alpha, 1, False)
full = ref.sim_matrix().copy()
ref.close()
lo, hi = multi.shard_ranges(n, world)[rank]
k = min(halo, n - hi)
sl, sd, se = l[lo:hi + k].clone(), d[lo:hi + k].clone(), e[lo:hi + k].clone()
per = (n_reads + world - 1) // world
r_lo, r_hi = min(rank * per, n_reads), min((rank + 1) * per, n_reads)
for sync_mode in (0, 1):
    ctx = LimeContext(local)
    ctx.set_stream(torch.cuda.current_stream().cuda_stream)
    h = torch.from_numpy(ctx.xchg_alloc(world, rank, per, n_refs, 1, 2_000_000, 1_500_000).copy()).to(dev)
    parts = [torch.empty_like(h) for _ in range(world)]
    dist.all_gather(parts, h)
    ctx.xchg_open(torch.stack(parts).cpu().numpy())
    dist.barrier()
    kk = ctx.xchg_counts_len()
    sent = torch.zeros(kk, dtype=torch.int64, device=dev)
    counts = torch.zeros(world * kk, dtype=torch.int64, device=dev)
    for rep in range(2):
        ctx.load(sl.data_ptr(), sd.data_ptr(), se.data_ptr(), n_reads, n_refs, n_own=hi - lo, n_halo=k, pos_base=lo,
                 is_last=rank == world - 1, mem=MEM_DEVICE, n_total=hi - lo + k)
        ctx.xchg_accumulate(alpha, 1, 0, sent.data_ptr())
        if sync_mode:
            torch.cuda.synchronize(); dist.barrier()
        dist.all_gather_into_tensor(counts, sent)
        if sync_mode:
            torch.cuda.synchronize(); dist.barrier()
        ctx.xchg_apply(0, counts.data_ptr())
        slab = ctx.sim_matrix(rows=per)
        st = ctx.xchg_status(0)
        exp = full[r_lo:r_hi]
        bad = int((slab[:r_hi - r_lo] != exp).sum())
        print(f"rank {rank} sync={sync_mode} rep={rep}: sent={sent.tolist()} counts={counts.view(world, kk).tolist()} "
              f"mismatching cells={bad} of {int((exp != 0).sum())} nonzero; slab sum {int(slab.astype(np.int64).sum())} "
              f"expected {int(exp.astype(np.int64).sum())}; status {st}", flush=True)
        if bad:
            rr, gg = np.nonzero(slab[:r_hi - r_lo] != exp)
            print(f"   rank {rank}: first bad cells (row, genome, got, want): "
                  f"{[(int(a), int(b), int(slab[a, b]), int(exp[a, b])) for a, b in list(zip(rr, gg))[:6]]}", flush=True)
        ctx.xchg_choose(0, 85, 0.25, r_hi - r_lo)          # zeroes the slab for the next repetition
        torch.cuda.synchronize(); dist.barrier()
    ctx.xchg_close()
    ctx.close()
dist.barrier()

# ---- part 2: the whole XchgJob (2 strands with DIFFERENT data) against the single-GPU job, slot by slot -------------
from lime_b200.pipeline import run_job, synthetic_lineage
rows_, v, m = synthetic_lineage(n_refs)
ctx = LimeContext(local)
ctx.set_stream(torch.cuda.current_stream().cuda_stream)
data = [(e, l, d)]
e2, l2, d2 = datagen.synth_arrays(n, n_reads, n_refs, 100, seed=78, device=dev, p_long=2e-5, max_seg=60000)
data.append((e2, l2, d2))
strands = [{"ebwt": a.data_ptr(), "lcp": b.data_ptr(), "da": c.data_ptr(), "n": n} for (a, b, c) in data]
exp, st1 = run_job(ctx, strands, n_reads, n_refs, 100, alpha, 0.25, 1, v, m, 1, True, mem=MEM_DEVICE, deferred=False)
exp = [np.array(x, copy=True) for x in exp]
sims = []
for (a, b, c) in data:
    ctx.load(b.data_ptr(), c.data_ptr(), a.data_ptr(), n_reads, n_refs, mem=MEM_DEVICE, n_total=n)
    ctx.cluster_similarity(alpha, 1, False)
    sims.append(ctx.sim_matrix().copy())
be = multi.CudaBackend(ctx, torch, dev)
shards = [dict(lcp=b[lo:hi + k].clone(), da=c[lo:hi + k].clone(), ebwt=a[lo:hi + k].clone(), n_own=hi - lo, n_halo=k, pos_base=lo,
               is_last=rank == world - 1) for (a, b, c) in data]
from oracle.oracle import Oracle          # diagnostic tool only
orc = Oracle()
for check_slabs in (1, 0):
    job = multi.XchgJob(dist, torch, be, rank, world, n_reads, n_refs, 2, 2_000_000, 1_500_000)
    for rep in range(2):
        job.begin()
        for s in range(2):
            if check_slabs:
                # run_strand by hand, looking at the slab between apply and choose
                be.load(shards[s]["lcp"], shards[s]["da"], shards[s]["ebwt"], shards[s]["n_own"], shards[s]["n_halo"], shards[s]["pos_base"],
                        shards[s]["is_last"], n_reads, n_refs)
                sent = be.xchg_accumulate(alpha, 1, s)
                counts = be.xchg_counts_buffer(s)
                dist.all_gather_into_tensor(counts, sent)
                be.xchg_apply(s, counts)
                slab = ctx.sim_matrix(rows=per)
                bad = int((slab[:r_hi - r_lo] != sims[s][r_lo:r_hi]).sum())
                print(f"rank {rank} rep={rep} strand {s}: slab mismatching cells={bad}; sent {sent.tolist()} counts {counts.tolist()}", flush=True)
                if bad:
                    rr, gg = np.nonzero(slab[:r_hi - r_lo] != sims[s][r_lo:r_hi])
                    print(f"   rank {rank}: bad cells (row, genome, got, want): {[(int(a), int(b), int(slab[a, b]), int(sims[s][r_lo + a, b])) for a, b in list(zip(rr, gg))[:10]]}; status {ctx.xchg_status(s)}", flush=True)
                be.xchg_choose(s, 85, 0.25, r_hi - r_lo)
                job._pending.append(s)
            else:
                job.run_strand(s, shards[s], alpha, 0.25, 1, 100)
        got = [np.array(x, copy=True) for x in job.finish(v, m, 1, True, gather=False)]
        diff = [int((np.asarray(g) != np.asarray(x[r_lo:r_hi])).sum()) for g, x in zip(got, exp)]
        print(f"rank {rank} check_slabs={check_slabs} rep={rep}: own-range assignment diffs (type, tax, sim) = {diff}", flush=True)
        for s in range(2):
            b, p = ctx.choose_result(s)
            bo, po = orc.choose(sims[s][r_lo:r_hi], 85, np.float32(0.25))
            print(f"   rank {rank} slot {s}: records {b.shape[0]} vs {bo.shape[0]}, bin equal {b.tobytes() == bo.tobytes()}, pos equal {p.tobytes() == po.tobytes()}", flush=True)
    job.close()
ctx.close()
dist.barrier()
dist.destroy_process_group()
