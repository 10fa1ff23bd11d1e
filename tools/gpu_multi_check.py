#!/usr/bin/env python
"""Multi-GPU parity check (run under torchrun, one rank per GPU, NCCL):
the position-sharded job (lime_b200.multi.ShardedJob over the C ABI) must give, bit for bit, the
assignments / nClusters / maxLen of the single-GPU job on the whole arrays.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 \
        --master-port 29511 tools/gpu_multi_check.py
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import torch
    import torch.distributed as dist
    from lime_b200 import datagen, multi
    from lime_b200.api import LimeContext, MEM_DEVICE
    from lime_b200.pipeline import run_job, synthetic_lineage

    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    ctx = LimeContext(local)
    ctx.set_stream(torch.cuda.current_stream().cuda_stream)

    ok = True
    big = os.environ.get("LIME_CHECK_BIG")          # one case of >= 50 M positions per strand (bench-scale kernels)
    cases = [(3_000_001, 30_000, 40, 1, 16, 2, 4096), (2_500_003, 20_000, 9, 0, 12, 4, 70_000), (2_000_003, 60_000, 930, 0, 16, 4, 70_000)]
    if big:
        cases.append((int(big), 300_000, 100, 1, 16, 2, 70_000))
    for (n, n_reads, n_refs, ebwt_mode, alpha, n_strands, halo) in cases:
        rows, v, m = synthetic_lineage(n_refs)
        whole, shards = [], []
        for s in range(n_strands):
            e, l, d = datagen.synth_arrays(n, n_reads, n_refs, 100, seed=77 + s, device=dev, p_long=2e-5, max_seg=60000)
            whole.append((e, l, d))
            lo, hi = multi.shard_ranges(n, world)[rank]
            own = hi - lo
            bufs = []
            for t in (l, d, e):
                b = torch.zeros(own + halo, dtype=t.dtype, device=dev)
                b[:own] = t[lo:hi]
                k = min(halo, n - hi)                       # halo taken from the whole array (same on all ranks)
                b[own:own + k] = t[hi:hi + k]
                bufs.append(b)
            shards.append(dict(lcp=bufs[0], da=bufs[1], ebwt=bufs[2], n_own=own, n_halo=min(halo, n - hi), pos_base=lo,
                               is_last=rank == world - 1))
        torch.cuda.synchronize()
        be = multi.CudaBackend(ctx, torch, dev)
        if os.environ.get("LIME_CHECK_TRIPLES"):
            raw_export = be.export

            def checked_export(lo, hi, buf, offset, _raw=raw_export, _nr=n_reads, _ng=n_refs):
                k = _raw(lo, hi, buf, offset)
                tr = buf[offset:offset + k].cpu().numpy().view(np.uint32).reshape(-1, 2)
                bad = (tr[:, 0] < lo) | (tr[:, 0] >= hi) | ((tr[:, 1] >> 8) >= _ng) | ((tr[:, 1] & 0xFF) == 0)
                if bad.any():
                    print(f"rank {rank}: export [{lo},{hi}) k={k} cap={buf.numel() - offset}: {int(bad.sum())} bad triples, first "
                          f"{tr[bad][:4].tolist()} at {np.nonzero(bad)[0][:4].tolist()} stats={ctx.sim_stats()}", flush=True)
                return k
            be.export = checked_export
            raw_import = be.import_

            def checked_import(buf, offset, k, _raw=raw_import, _nr=n_reads, _ng=n_refs):
                tr = buf[offset:offset + k].cpu().numpy().view(np.uint32).reshape(-1, 2)
                bad = (tr[:, 0] >= _nr) | ((tr[:, 1] >> 8) >= _ng)
                print(f"rank {rank}: import k={k} bad={int(bad.sum())} first bad {tr[bad][:4].tolist()} at {np.nonzero(bad)[0][:4].tolist()} "
                      f"ctx n_reads={ctx.n_reads} n_refs={ctx.n_refs}", flush=True)
                return _raw(buf, offset, k)
            be.import_ = checked_import
        job = multi.ShardedJob(dist, torch, be, rank, world, n_reads, n_refs)
        got = None
        if not os.environ.get("LIME_CHECK_SKIP_EXACT"):
            for s, sh in enumerate(shards):
                job.run_strand(s, sh, alpha, 0.25, ebwt_mode, 100)
            got = job.finish(v, m, 1, True)
        # the same job with the rows on the read owner and the updates routed as records over NVLink (XchgJob)
        flat_cap, inbox_cap = multi.xchg_capacities(max(sh["n_own"] + sh["n_halo"] for sh in shards), world, density=1.0)
        pj = multi.XchgJob(dist, torch, be, rank, world, n_reads, n_refs, n_strands, flat_cap, inbox_cap)
        got_p, carried = None, False
        try:
            for rep in range(2):                       # twice: slabs and inboxes are reused
                pj.begin()
                for s, sh in enumerate(shards):
                    pj.run_strand(s, sh, alpha, 0.25, ebwt_mode, 100)
                got_p = [np.array(a) for a in pj.finish(v, m, 1, True)]
                xst = [be.xchg_status(s) for s in range(n_strands)]
        except multi.CarryDetected as ex:
            # legitimate only if the single-GPU job takes the exact path too (checked below)
            carried, got_p = True, None
            print(f"rank {rank}: record exchange refused the job: {ex}", flush=True)
        pj.close()
        # single-GPU job on the whole arrays (every rank computes it, cheap at this size)
        strands = [{"ebwt": e.data_ptr(), "lcp": l.data_ptr(), "da": d.data_ptr(), "n": n} for (e, l, d) in whole]
        exp, stats = run_job(ctx, strands, n_reads, n_refs, 100, alpha, 0.25, ebwt_mode, v, m, 1, True, mem=MEM_DEVICE)
        if got is None:
            got, job.stats = exp, [{"n_clusters": st[0], "max_len": st[1], "triples_sent": 0} for st in stats]
        same = all(np.array_equal(np.asarray(a).view(np.uint8), np.asarray(b).view(np.uint8)) for a, b in zip(got, exp))
        if carried:
            # a cell above 255 must also show on one GPU: the synchronising calls then take the exact path
            exact = 0
            for (e_, l_, d_) in whole:
                ctx.load(l_.data_ptr(), d_.data_ptr(), e_.data_ptr() if ebwt_mode else None, n_reads, n_refs, mem=MEM_DEVICE, n_total=n)
                ctx.cluster_similarity(alpha, ebwt_mode, False)
                exact += int(ctx.sim_stats()["exact_pass"])
            same_p = exact > 0
            got_p = exp
        else:
            same_p = all(np.array_equal(np.asarray(a).view(np.uint8), np.asarray(b).view(np.uint8)) for a, b in zip(got_p, exp))
        sc = all(js["n_clusters"] == st[0] and js["max_len"] == st[1] for js, st in zip(job.stats, stats))
        sc = sc and (carried or all(js["n_clusters"] == st[0] and js["max_len"] == st[1] for js, st in zip(pj.stats, stats)))
        if not same_p:
            for s in range(n_strands):
                e_, l_, d_ = whole[s]
                ctx.load(l_.data_ptr(), d_.data_ptr(), e_.data_ptr() if ebwt_mode else None, n_reads, n_refs, mem=MEM_DEVICE, n_total=n)
                ctx.cluster_similarity(alpha, ebwt_mode, False)
                simm = ctx.sim_matrix()
                r_lo, r_hi = pj.lo, pj.hi
                print(f"rank {rank}: strand {s}: expected sum of own rows {int(simm[r_lo:r_hi].astype(np.int64).sum())}, nonzero cells {int((simm[r_lo:r_hi] != 0).sum())}; "
                      f"xchg status of the last job: {xst[s]}", flush=True)
            for name, a, b in zip(("type", "tax", "sim"), got_p, exp):
                bad = np.flatnonzero(np.asarray(a) != np.asarray(b))
                print(f"rank {rank}: record-exchange {name}: {bad.shape[0]} of {len(b)} reads differ, first {bad[:6].tolist()} "
                      f"got {np.asarray(a)[bad[:6]].tolist()} want {np.asarray(b)[bad[:6]].tolist()}; xchg stats {pj.stats}", flush=True)
        kinds = {chr(t): int((got[0] == t).sum()) for t in np.unique(got[0])}
        print(f"rank {rank}/{world}: n={n} strands={n_strands} ebwt={ebwt_mode} alpha={alpha}: assignments equal={same} record-exchange job equal={same_p} "
              f"scalars equal={sc} {kinds} triples sent={[js['triples_sent'] for js in job.stats]}", flush=True)
        ok = ok and same and same_p and sc
    flag = torch.tensor([0 if ok else 1], device=dev)
    dist.all_reduce(flag)
    dist.barrier()
    ctx.close()
    dist.destroy_process_group()
    if int(flag.item()):
        raise SystemExit("MULTI-GPU PARITY FAILED")
    if rank == 0:
        print("multi-GPU parity ok")


if __name__ == "__main__":
    main()
