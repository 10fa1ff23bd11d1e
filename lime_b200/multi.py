"""Position-sharded multi-GPU driver of the LiME hot path (one process per GPU, torch.distributed).

Sharding (north_star):
  * every strand's position range is cut into contiguous shards, one per rank; a cluster belongs
    to the shard that holds its head -- the thread-chunk rule of ClusterLCP.cpp:150-202 -- so each
    rank also needs the first `halo` positions of the next shard (the "run past the chunk end" of
    :246-264).  `exchange_halo` moves them rank k+1 -> rank k (point to point over NVLink);
  * reads are owned by contiguous ranges and the similarity rows of a read live on its owner only.
    `XchgJob` (the default): every update (read, genome, t) of a rank's positions is appended to a local
    list by the fused kernel, sorted by owner and written into the owner's inbox over NVLink
    (lime_xchg_accumulate, bulk stores into CUDA-IPC mapped peer memory); one tiny all-gather of the
    record counts per strand is the only collective and the barrier; the owner adds the records to
    its rows (lime_xchg_apply).  Nothing synchronises with the host until Classify;
  * `ShardedJob` (the exact route, used when a cell exceeds 255 and must wrap like the reference's
    uint8 cell, Tools.h:68): each rank accumulates a PARTIAL matrix over all reads, the non-zero
    cells of the other ranks' ranges leave as (read, genome, count) triples (lime_sim_export), are
    routed by read owner with ONE all-to-all (NCCL) and added on arrival (lime_sim_import);
  * the owner runs clusterChoose and, after all strands, Classify on its read range; the
    assignments are gathered.
nClusters / maxLen are the reference's reduction(+) / reduction(max) (ClusterLCP.cpp:140): one
all-reduce each.

The compute goes through a small backend object so that the same host logic runs on CUDA
(`CudaBackend`, the C ABI) and, in the CPU tests (gloo, world size 2), on a checker backend.
"""
import os

import numpy as np

from .pipeline import norm_of


def read_ranges(n_reads, world):
    """Contiguous read ownership: rank k owns [lo_k, hi_k)."""
    per = (n_reads + world - 1) // world
    return [(min(k * per, n_reads), min((k + 1) * per, n_reads)) for k in range(world)]


def shard_ranges(n, world):
    """Contiguous position shards of an n-position strand."""
    per = (n + world - 1) // world
    return [(min(k * per, n), min((k + 1) * per, n)) for k in range(world)]


def exchange_halo(dist, rank, world, tensors, n_own, halo):
    """tensors: 1-D torch tensors with room for n_own + halo elements.  The first `halo` owned
    elements go to rank-1, the tail [n_own, n_own+halo) is filled from rank+1.  Returns the number
    of halo elements this rank now holds (0 on the last rank)."""
    reqs = []
    for t in tensors:
        if t is None:
            continue
        if rank > 0:
            reqs.append(dist.isend(t[:halo].contiguous(), rank - 1))
        if rank < world - 1:
            reqs.append(dist.irecv(t[n_own:n_own + halo], rank + 1))
    for r in reqs:
        r.wait()
    return halo if rank < world - 1 else 0


class CudaBackend:
    """The C ABI (LimeContext) behind the backend interface; exchange buffers are CUDA tensors."""

    def __init__(self, ctx, torch, device):
        self.ctx, self.torch, self.device = ctx, torch, device

    def load(self, lcp, da, ebwt, n_own, n_halo, pos_base, is_last, n_reads, n_refs):
        """CUDA tensors are borrowed (LIME_MEM_DEVICE); CPU tensors (page-locked for full PCIe rate) are
        copied in by lime_load(LIME_MEM_HOST) -- the end-to-end leg of the bench."""
        from ._lib import MEM_DEVICE, MEM_HOST
        self._keep = (lcp, da, ebwt)
        n = n_own + n_halo
        if lcp.device.type == "cpu":
            self.ctx.load(lcp[:n].numpy().view(np.uint32), da[:n].numpy().view(np.uint32),
                          ebwt[:n].numpy() if ebwt is not None else None, n_reads, n_refs,
                          n_own=n_own, n_halo=n_halo, pos_base=pos_base, is_last=is_last, mem=MEM_HOST)
            return
        self.ctx.load(lcp.data_ptr(), da.data_ptr(), ebwt.data_ptr() if ebwt is not None else None, n_reads, n_refs,
                      n_own=n_own, n_halo=n_halo, pos_base=pos_base, is_last=is_last, mem=MEM_DEVICE,
                      n_total=n)

    def cluster_similarity(self, alpha, ebwt_mode):
        return self.ctx.cluster_similarity(alpha, ebwt_mode, False)

    def updates(self):
        return int(self.ctx.sim_stats()["updates"])

    def new_buffer(self, n):
        return self.torch.empty(max(int(n), 1), dtype=self.torch.int64, device=self.device)

    def export(self, lo, hi, buf, offset):
        return self.ctx.sim_export(lo, hi, buf.data_ptr() + 8 * offset, buf.numel() - offset)

    def import_(self, buf, offset, n):
        self.ctx.sim_import(buf.data_ptr() + 8 * offset, n)

    def choose(self, slot, norm, beta, lo, hi):
        return self.ctx.choose(slot, norm, beta, lo, hi)

    def classify(self, n_files, n_refs, v, m, tax_rank, higher):
        return self.ctx.classify(n_files, n_refs, v, m, tax_rank, higher)

    # ---- rows on the read owner, updates routed as records (XchgJob) ---------------------------------
    def xchg_setup(self, dist, world, rank, per, n_refs, n_strands, flat_cap, inbox_cap):
        """Allocate slabs / update list / inbox, all-gather the CUDA IPC handles, map every rank's inbox."""
        torch = self.torch
        self.ctx.set_stream(torch.cuda.current_stream().cuda_stream)
        mine = torch.from_numpy(self.ctx.xchg_alloc(world, rank, per, n_refs, n_strands, flat_cap, inbox_cap).copy()).to(self.device)
        parts = [torch.empty_like(mine) for _ in range(world)]
        dist.all_gather(parts, mine)
        self.ctx.xchg_open(torch.stack(parts).cpu().numpy() if world > 1 else None)
        dist.barrier()                       # every rank has mapped every inbox (and cleared its mailbox)

    def xchg_accumulate(self, alpha, ebwt_mode, strand):
        self.ctx.xchg_accumulate(alpha, ebwt_mode, strand)

    def xchg_apply(self, strand):
        self.ctx.xchg_apply(strand)

    def xchg_choose(self, strand, norm, beta, n_own):
        self.ctx.xchg_choose(strand, norm, beta, n_own)

    def xchg_status(self, strand):
        return self.ctx.xchg_status(strand)

    def strand_status(self, slot):
        return self.ctx.strand_status(slot)

    def xchg_close(self):
        self.ctx.xchg_close()

    def classify_tensors(self, n_files, n_refs, v, m, tax_rank, higher, pad_to):
        """Classify and leave the assignments on the device: (type u8, taxid i32, sim-bits i32) tensors
        of pad_to elements (zero beyond this rank's reads), ready for an all-gather over NVLink."""
        torch = self.torch
        k = self.ctx.classify_on_device(n_files, n_refs, v, m, tax_rank, higher)
        typ = torch.zeros(pad_to, dtype=torch.uint8, device=self.device)
        tax = torch.zeros(pad_to, dtype=torch.int32, device=self.device)
        sim = torch.zeros(pad_to, dtype=torch.int32, device=self.device)
        if k:
            self.ctx.classify_result_to_device(typ.data_ptr(), tax.data_ptr(), sim.data_ptr())
        return typ, tax, sim


class ShardedJob:
    """One classification job (2 or 4 strands) over `world` ranks.

        job = ShardedJob(dist, torch, backend, rank, world, n_reads, n_refs)
        for s, shard in enumerate(strands):           # this rank's shard of every strand
            job.run_strand(s, shard, alpha, beta, ebwt_mode, read_len)
        typ, tax, sim = job.finish(v, m, tax_rank, higher)      # all reads, on every rank
    """

    def __init__(self, dist, torch, backend, rank, world, n_reads, n_refs, timer=None):
        self.dist, self.torch, self.be = dist, torch, backend
        self.rank, self.world = rank, world
        self.n_reads, self.n_refs = n_reads, n_refs
        self.ranges = read_ranges(n_reads, world)
        self.lo, self.hi = self.ranges[rank]
        self.timer = timer or (lambda name: _NoTimer())
        self.sendbuf = self.recvbuf = None
        self._pinned = {}
        self.stats = []
        self.n_strands = 0

    def _cpu_or_dev(self, values):
        t = self.torch.tensor(values, dtype=self.torch.int64)
        dev = getattr(self.be, "device", None)
        return t.to(dev) if dev is not None else t

    def run_strand(self, slot, shard, alpha, beta, ebwt_mode, read_len):
        """shard: dict(lcp, da, ebwt, n_own, n_halo, pos_base, is_last) -- halo already in place."""
        be, dist, torch = self.be, self.dist, self.torch
        with self.timer("compute"):
            be.load(shard["lcp"], shard["da"], shard["ebwt"] if ebwt_mode else None, shard["n_own"], shard["n_halo"],
                    shard["pos_base"], shard["is_last"], self.n_reads, self.n_refs)
            n_cl, max_len = be.cluster_similarity(alpha, ebwt_mode)
        # ---- export: non-zero cells of the other owners' read ranges, in destination order -----------
        with self.timer("export"):
            need = be.updates()                      # every non-zero cell received >= 1 update
            if self.sendbuf is None or self.sendbuf.numel() < need:
                self.sendbuf = be.new_buffer(need)
            send_counts, off = [], 0
            for dst, (lo, hi) in enumerate(self.ranges):
                n = 0
                if dst != self.rank and hi > lo:
                    n = be.export(lo, hi, self.sendbuf, off)
                    assert off + n <= self.sendbuf.numel(), "triple bound violated"
                send_counts.append(n)
                off += n
        # ---- route by read owner: counts, then ONE all-to-all of the triples --------------------------
        with self.timer("counts"):
            # every rank learns the whole count matrix with one all-gather (an equal-split
            # all_to_all_single of a few integers was measured at 32 ms per call under NCCL 2.28)
            sc = self._cpu_or_dev(send_counts)
            rows = [torch.empty_like(sc) for _ in range(self.world)]
            dist.all_gather(rows, sc)
            recv_counts = [int(r[self.rank].item()) for r in rows]
            total_in = sum(recv_counts)
            if self.recvbuf is None or self.recvbuf.numel() < max(total_in, 1):
                self.recvbuf = be.new_buffer(total_in)
        with self.timer("all_to_all"):
            dist.all_to_all_single(self.recvbuf[:total_in], self.sendbuf[:off], recv_counts, send_counts)
        # ---- merge on the owner: one source at a time (keys are unique within a source) --------------
        with self.timer("import"):
            o = 0
            for src, n in enumerate(recv_counts):
                if n:
                    be.import_(self.recvbuf, o, n)
                o += n
        with self.timer("choose"):
            n_rec = be.choose(slot, norm_of(read_len, alpha), beta, self.lo, self.hi)
        # ---- scalars of <stem>.out: reduction(+:nClusters) reduction(max:maxLen), ClusterLCP.cpp:140 ---
        with self.timer("scalars"):
            s = self._cpu_or_dev([n_cl])
            mx = self._cpu_or_dev([max_len])
            dist.all_reduce(s, op=dist.ReduceOp.SUM)
            dist.all_reduce(mx, op=dist.ReduceOp.MAX)
        self.stats.append({"n_clusters": int(s.item()), "max_len": int(mx.item()), "n_records_local": int(n_rec),
                           "triples_sent": int(off), "triples_received": int(total_in)})
        self.n_strands = max(self.n_strands, slot + 1)
        return self.stats[-1]

    def classify_own(self, v, m=None, tax_rank=1, higher=False, gather=True):
        """Classify on the owned read range -> ("dev", device tensors padded for the gather) or ("host", arrays)."""
        per = self.ranges[0][1] - self.ranges[0][0]
        on_device = hasattr(self.be, "classify_tensors") and gather and self.world > 1
        with self.timer("classify"):
            if on_device:
                return "dev", self.be.classify_tensors(self.n_strands, self.n_refs, v, m, tax_rank, higher, per)
            return "host", self.be.classify(self.n_strands, self.n_refs, v, m, tax_rank, higher)

    def finish(self, v, m=None, tax_rank=1, higher=False, gather=True):
        """Classify on the owned read range, then gather the assignments of all reads.
        gather=True: every rank returns them; "root": rank 0 returns them, the others (None,)*3;
        False: every rank returns its own read range only."""
        res = self.classify_own(v, m, tax_rank, higher, gather)
        if not gather or self.world == 1:
            return res[1]          # the backend's (page-locked, context-owned) arrays: valid until the next classify
        return self.gather_assignments(res, gather)

    def gather_assignments(self, res, gather=True):
        torch, dist = self.torch, self.dist
        per = self.ranges[0][1] - self.ranges[0][0]
        k = self.hi - self.lo
        on_device = res[0] == "dev"
        with self.timer("gather"):
            # three contiguous all-gathers (type u8, taxid u32, sim f32 bits); every rank owns `per`
            # reads except possibly the last ones, so the padded concatenation IS the read order
            if on_device:
                mine = res[1]
            else:
                typ, tax, sim = res[1]
                mine = []
                for arr, tdt, ndt in ((typ, torch.uint8, np.uint8), (tax, torch.int32, np.int32), (sim, torch.int32, np.int32)):
                    t = torch.zeros(per, dtype=tdt)
                    t[:k] = torch.from_numpy(np.ascontiguousarray(arr).view(ndt)[:k].copy())
                    dev = getattr(self.be, "device", None)
                    mine.append(t.to(dev) if dev is not None else t)
            outs = []
            fetch = gather is True or self.rank == 0
            for t in mine:
                parts = [torch.empty_like(t) for _ in range(self.world)]
                dist.all_gather(parts, t)
                if fetch:
                    whole = torch.cat(parts)[:self.n_reads]
                    if whole.is_cuda:
                        key = (len(outs), whole.shape[0], whole.dtype)
                        host = self._pinned.get(key)
                        if host is None:              # page-locked once, reused by every later job step
                            host = self._pinned[key] = torch.empty(whole.shape, dtype=whole.dtype, pin_memory=True)
                        host.copy_(whole, non_blocking=True)
                        outs.append(host)
                    else:
                        outs.append(whole)
            if fetch and mine and mine[0].is_cuda:
                torch.cuda.synchronize()
        if not fetch:
            return None, None, None
        # views of job-owned (page-locked) buffers on CUDA: valid until the next finish()
        return outs[0].numpy(), outs[1].numpy().view(np.uint32), outs[2].numpy().view(np.float32)


class CarryDetected(RuntimeError):
    """The exchange job cannot stand: a similarity cell exceeded 255 (redo it with ShardedJob, whose
    backend reruns the pass into exact 32-bit cells -- the uint8 wrap of SimArray_, Tools.h:68), or a
    record list / inbox was too small (`overflow`: redo it with larger lists)."""

    def __init__(self, msg, overflow=False):
        super().__init__(msg)
        self.overflow = overflow


def xchg_capacities(n_positions, world, density=0.25):
    """Record capacities of an XchgJob for shards of n_positions: the update list of a strand (the bench
    workload issues 0.11 updates per position; every warp that appends takes whole blocks of 256 slots, hence
    the 3 Mi of slack) and one inbox region (a source's share for one owner)."""
    flat_cap = int(n_positions * density) + (3 << 20)
    inbox_cap = int(flat_cap * 1.5 / world) + (1 << 16)
    return flat_cap, inbox_cap


class XchgJob:
    """The job of ShardedJob with the similarity rows on the read owner only: compute and routing are
    queued per strand without a host synchronisation or a collective (lime_xchg_accumulate -> lime_xchg_apply
    -> lime_xchg_choose; records, counts and "delivered" flags are stores to peer memory, and the routing /
    applying / choosing of a strand runs beside the scoring of the next one); Classify reads the strands'
    state back, then ONE all-gather lets the ranks agree on the outcome and the scalars -- it is also the
    meeting point the next job needs before it may rewrite the inboxes.

        job = XchgJob(dist, torch, backend, rank, world, n_reads, n_refs, n_strands, flat_cap, inbox_cap)
        for s, shard in enumerate(strands): job.run_strand(s, shard, alpha, beta, ebwt_mode, read_len)
        typ, tax, sim = job.finish(v, m, tax_rank, higher)
    """

    def __init__(self, dist, torch, backend, rank, world, n_reads, n_refs, n_strands, flat_cap, inbox_cap, timer=None):
        self.dist, self.torch, self.be = dist, torch, backend
        self.rank, self.world = rank, world
        self.n_reads, self.n_refs, self.n_strands = n_reads, n_refs, n_strands
        self.ranges = read_ranges(n_reads, world)
        self.lo, self.hi = self.ranges[rank]
        self.per = max(self.ranges[0][1] - self.ranges[0][0], 1)
        self.timer = timer or (lambda name: _NoTimer())
        self.stats, self._pending = [], []
        self._inner = ShardedJob(dist, torch, backend, rank, world, n_reads, n_refs)   # finish() = classify + gather
        # a source addresses its region inside the owner's inbox with the region size: the capacities must be
        # the same on every rank (shards differ in length, e.g. the last one has no halo) -> agree on the maximum
        caps = self._inner._cpu_or_dev([int(flat_cap), int(inbox_cap)])
        dist.all_reduce(caps, op=dist.ReduceOp.MAX)
        self.flat_cap, self.inbox_cap = int(caps[0].item()), int(caps[1].item())
        backend.xchg_setup(dist, world, rank, self.per, n_refs, n_strands, self.flat_cap, self.inbox_cap)
        self.closed = False

    def begin(self):
        if self.closed:
            raise RuntimeError("XchgJob is closed")
        self.stats, self._pending = [], []

    def run_strand(self, slot, shard, alpha, beta, ebwt_mode, read_len):
        be, dist = self.be, self.dist
        with self.timer("compute"):
            be.load(shard["lcp"], shard["da"], shard["ebwt"] if ebwt_mode else None, shard["n_own"], shard["n_halo"],
                    shard["pos_base"], shard["is_last"], self.n_reads, self.n_refs)
            be.xchg_accumulate(alpha, ebwt_mode, slot)          # + routing, queued on the library's side stream
        with self.timer("apply"):
            be.xchg_apply(slot)                                 # side stream: waits for every source's flag on the device
        with self.timer("choose"):
            be.xchg_choose(slot, norm_of(read_len, alpha), beta, self.hi - self.lo)
        self._pending.append(slot)

    def finish(self, v, m=None, tax_rank=1, higher=False, gather=True):
        from ._lib import ERR_RETRY, LimeError
        torch, dist = self.torch, self.dist
        inner = self._inner
        inner.timer, inner.n_strands = self.timer, len(self._pending)
        code, res, msg = 0, None, ""
        try:
            res = inner.classify_own(v, m, tax_rank, higher, gather)      # also reads the strands' state back
        except LimeError as e:
            code, msg = e.code, str(e)
        with self.timer("check"):
            # the ranks agree on the outcome (every rank takes the same branch) and on the scalars: one all-gather
            st = [self.be.xchg_status(s) for s in self._pending]
            over = int(any(x["list_slots"] > self.flat_cap or (x["err"] & 4) for x in st))
            k = len(st)
            mine = inner._cpu_or_dev([code, over] + [x["n_clusters"] for x in st] + [x["updates"] for x in st]
                                     + [x["max_len"] for x in st])
            rows = torch.empty(self.world * mine.numel(), dtype=mine.dtype, device=mine.device)
            dist.all_gather_into_tensor(rows, mine)
            rows = rows.cpu().view(self.world, -1)
            code_all, over_all = int(rows[:, 0].min().item()), int(rows[:, 1].max().item())
            if code_all == ERR_RETRY:
                raise CarryDetected(msg or "another rank must redo its strands", overflow=bool(over_all))
            if code_all != 0:
                raise LimeError(code_all, msg or "a rank of the job failed")
            tot, mx = rows[:, 2:2 + 2 * k].sum(dim=0), rows[:, 2 + 2 * k:].max(dim=0).values
            self.stats = [{"n_clusters": int(tot[i].item()), "max_len": int(mx[i].item()), "updates": int(tot[k + i].item()),
                           "n_records_local": st[i]["n_records"]} for i in range(k)]
        if not gather or self.world == 1:
            return res[1]
        return inner.gather_assignments(res, gather)

    def close(self):
        """Collective: unmaps and frees slabs and inboxes; the backend then allocates whole matrices again."""
        if self.closed:
            return
        self.closed = True
        self.dist.barrier()
        self.be.xchg_close()


def run_sharded(dist, torch, backend, rank, world, n_reads, n_refs, shards, alpha, beta, ebwt_mode, read_len,
                v, m=None, tax_rank=1, higher=False, gather=True, xchg_job=None):
    """One classification job over `world` ranks: through `xchg_job` (an XchgJob, reused between calls)
    when given, and again through the exact exchange route when that one reports a cell above 255 (or lists
    that were too small) -- every rank sees the same all-reduced outcome, so all ranks take the same branch;
    the exchange job is closed then (its buffers are released so that the backend allocates whole matrices
    again).  shards: this rank's shard of every strand (dicts as for run_strand).
    Returns (assignments, per-strand stats, route)."""
    if xchg_job is not None and not xchg_job.closed:
        try:
            xchg_job.begin()
            for s, sh in enumerate(shards):
                xchg_job.run_strand(s, sh, alpha, beta, ebwt_mode, read_len)
            return xchg_job.finish(v, m, tax_rank, higher, gather=gather), list(xchg_job.stats), "xchg"
        except CarryDetected:
            xchg_job.close()
    job = ShardedJob(dist, torch, backend, rank, world, n_reads, n_refs)
    for s, sh in enumerate(shards):
        job.run_strand(s, sh, alpha, beta, ebwt_mode, read_len)
    return job.finish(v, m, tax_rank, higher, gather=gather), list(job.stats), "exchange"


def run_jobs_in_flight(jobs, k_steps, shards, alpha, beta, ebwt_mode, read_len, v, m=None, tax_rank=1, higher=False, gather=False):
    """k_steps classification jobs over the same shards with len(jobs) of them in flight: `jobs` are XchgJobs on
    different backends (contexts) of this rank; the strands of job k + 1 are queued -- run_strand never waits on a
    CUDA backend -- before finish() of job k is called, so job k's routing / applying / clusterChoose tail, its
    Classify, result copy and the ranks' agreement run beside the scoring of job k + 1.  Every rank must call this with
    the same k_steps (finish() holds a collective).  Returns the assignments of the last job; CarryDetected from a
    finish() propagates after the jobs already queued have been drained (their results are dropped)."""
    assert jobs and k_steps >= 1
    pending, out = [], None
    try:
        for k in range(k_steps):
            j = jobs[k % len(jobs)]
            if len(pending) == len(jobs):                   # this context's previous job first
                out = pending.pop(0).finish(v, m, tax_rank, higher, gather=gather)
            j.begin()
            for s, sh in enumerate(shards):
                j.run_strand(s, sh, alpha, beta, ebwt_mode, read_len)
            pending.append(j)
            if len(pending) > 1 or len(jobs) == 1:
                out = pending.pop(0).finish(v, m, tax_rank, higher, gather=gather)
        while pending:
            out = pending.pop(0).finish(v, m, tax_rank, higher, gather=gather)
        return out
    except CarryDetected:
        while pending:                                      # every rank drains the same jobs: the collectives stay matched
            try:
                pending.pop(0).finish(v, m, tax_rank, higher, gather=gather)
            except CarryDetected:
                pass
        raise


class _NoTimer:
    def __enter__(self):
        return self

    def __exit__(self, *a):
        return False
