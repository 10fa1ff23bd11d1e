"""Position-sharded multi-GPU driver of the LiME hot path (one process per GPU, torch.distributed).

Sharding (north_star):
  * every strand's position range is cut into contiguous shards, one per rank; a cluster belongs
    to the shard that holds its head -- the thread-chunk rule of ClusterLCP.cpp:150-202 -- so each
    rank also needs the first `halo` positions of the next shard (the "run past the chunk end" of
    :246-264).  `exchange_halo` moves them rank k+1 -> rank k (point to point over NVLink);
  * each rank accumulates a PARTIAL similarity matrix over all reads from its positions
    (lime_cluster_similarity);
  * reads are owned by contiguous ranges; the non-zero cells of the other ranks' ranges leave as
    (read, genome, count) triples (lime_sim_export), are routed by read owner with ONE all-to-all
    (NCCL over NVLink / NVSwitch) and added on arrival (lime_sim_import, uint8 wrap = the
    reference's cell, so sum-then-wrap equals the single-GPU result);
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

    # ---- peer-memory accumulation (PeerJob) -------------------------------------------------------
    def peer_setup(self, dist, world, rank, rows, n_refs, n_slabs):
        """Allocate this rank's slabs, all-gather the CUDA IPC handles, map every owner's slabs."""
        torch = self.torch
        mine = torch.from_numpy(self.ctx.peer_alloc(rows, n_refs, n_slabs).copy()).to(self.device)
        parts = [torch.empty_like(mine) for _ in range(world)]
        dist.all_gather(parts, mine)
        self.ctx.peer_open(world, rank, torch.stack(parts).cpu().numpy(), rows)
        dist.barrier()                       # every rank has mapped every slab

    def peer_zero(self):
        self.ctx.peer_zero()

    def peer_accumulate(self, alpha, ebwt_mode, slab):
        return self.ctx.peer_accumulate(alpha, ebwt_mode, slab)

    def peer_select(self, slab, total_updates, byte_sums, k):
        """byte_sums: int64 tensor of the job; element k receives the slab's byte sum (asynchronously)."""
        self.ctx.peer_select(slab, total_updates, byte_sums.data_ptr() + 8 * k)

    def peer_zero_slab(self, slab):
        self.ctx.peer_zero_slab(slab)

    def peer_close(self):
        self.ctx.peer_close()

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

    def finish(self, v, m=None, tax_rank=1, higher=False, gather=True):
        """Classify on the owned read range, then gather the assignments of all reads.
        gather=True: every rank returns them; "root": rank 0 returns them, the others (None,)*3;
        False: every rank returns its own read range only."""
        torch, dist = self.torch, self.dist
        per = self.ranges[0][1] - self.ranges[0][0]
        k = self.hi - self.lo
        on_device = hasattr(self.be, "classify_tensors") and gather and self.world > 1
        with self.timer("classify"):
            if on_device:
                mine = self.be.classify_tensors(self.n_strands, self.n_refs, v, m, tax_rank, higher, per)
            else:
                typ, tax, sim = self.be.classify(self.n_strands, self.n_refs, v, m, tax_rank, higher)
        if not gather or self.world == 1:
            return typ, tax, sim          # the backend's (page-locked, context-owned) arrays: valid until the next classify
        with self.timer("gather"):
            # three contiguous all-gathers (type u8, taxid u32, sim f32 bits); every rank owns `per`
            # reads except possibly the last ones, so the padded concatenation IS the read order
            if not on_device:
                mine = []
                for arr, tdt, ndt in ((typ, torch.uint8, np.uint8), (tax, torch.int32, np.int32), (sim, torch.int32, np.int32)):
                    t = torch.zeros(per, dtype=tdt)
                    t[:k] = torch.from_numpy(np.ascontiguousarray(arr).view(ndt)[:k].copy())
                    mine.append(t)
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
            if fetch and on_device:
                torch.cuda.synchronize()
        if not fetch:
            return None, None, None
        # views of job-owned (page-locked) buffers on CUDA: valid until the next finish()
        return outs[0].numpy(), outs[1].numpy().view(np.uint32), outs[2].numpy().view(np.float32)


class CarryDetected(RuntimeError):
    """A similarity cell exceeded 255 during a peer-memory job: redo it with ShardedJob, whose
    backend reruns the pass into exact 32-bit cells (the uint8 wrap of SimArray_, Tools.h:68)."""


class PeerJob:
    """The same job as ShardedJob with the exchange folded into the compute kernel: every rank holds
    only the similarity rows of the reads it owns (one slab per strand) and maps the slabs of all other
    ranks (CUDA IPC over NVLink / NVSwitch); the fused kernel adds every (read, genome, t) straight
    into the owner's slab (lime_peer_accumulate).  No partial matrices, no triples, no all-to-all:
    the collectives left are three tiny all-gathers per job that double as the barriers
      zeroed slabs -> [adds of all strands] -> scalars -> [choose, classify] -> byte sums (carry check).

        job = PeerJob(dist, torch, backend, rank, world, n_reads, n_refs, n_strands)
        job.begin()
        for s, shard in enumerate(strands): job.run_strand(s, shard, alpha, beta, ebwt_mode, read_len)
        typ, tax, sim = job.finish(v, m, tax_rank, higher)
    """

    def __init__(self, dist, torch, backend, rank, world, n_reads, n_refs, n_strands, timer=None):
        self.dist, self.torch, self.be = dist, torch, backend
        self.rank, self.world = rank, world
        self.n_reads, self.n_refs, self.n_strands = n_reads, n_refs, n_strands
        self.ranges = read_ranges(n_reads, world)
        self.lo, self.hi = self.ranges[rank]
        self.per = max(self.ranges[0][1] - self.ranges[0][0], 1)
        self.timer = timer or (lambda name: _NoTimer())
        self.stats, self._local, self._pending = [], [], []
        self._inner = ShardedJob(dist, torch, backend, rank, world, n_reads, n_refs)   # finish() = classify + gather
        backend.peer_setup(dist, world, rank, self.per, n_refs, n_strands)
        self._dirty, self.closed = False, False
        self._bs = torch.zeros(n_strands, dtype=torch.int64)
        if getattr(backend, "device", None) is not None:
            self._bs = self._bs.to(backend.device)

    def _all_gather_i64(self, values):
        """Every rank's vector of integers, as a (world, len) int64 numpy array; also a barrier."""
        torch, dist = self.torch, self.dist
        t = self._inner._cpu_or_dev(values)
        parts = [torch.empty_like(t) for _ in range(self.world)]
        dist.all_gather(parts, t)
        return torch.stack(parts).cpu().numpy().reshape(self.world, -1)

    def begin(self):
        """Nobody may add before every rank's slabs are zero: true after the constructor (fresh slabs,
        then the handle all-gather) and after every finish() (slabs cleared before its last all-gather);
        an abandoned job needs the explicit zero + barrier."""
        if self.closed:
            raise RuntimeError("PeerJob is closed")
        if self._dirty:
            with self.timer("zero"):
                self.be.peer_zero()
                self.dist.barrier()
        self._dirty = True
        self.stats, self._local, self._pending = [], [], []

    def run_strand(self, slot, shard, alpha, beta, ebwt_mode, read_len):
        be = self.be
        with self.timer("compute"):
            be.load(shard["lcp"], shard["da"], shard["ebwt"] if ebwt_mode else None, shard["n_own"], shard["n_halo"],
                    shard["pos_base"], shard["is_last"], self.n_reads, self.n_refs)
            n_cl, max_len, updates, sum_t = be.peer_accumulate(alpha, ebwt_mode, slot)
        self._local += [n_cl, max_len, updates, sum_t]
        self._pending.append((slot, norm_of(read_len, alpha), beta))

    def finish(self, v, m=None, tax_rank=1, higher=False, gather=True):
        be = self.be
        with self.timer("scalars"):
            # every rank's (nClusters, maxLen, updates, sum_t) per strand; completing it means every
            # rank's adds have landed (each rank synchronised its kernels before contributing)
            sc = self._all_gather_i64(self._local).reshape(self.world, len(self._pending), 4)
        with self.timer("choose"):
            for k, (slot, norm, beta) in enumerate(self._pending):
                be.peer_select(slot, int(sc[:, k, 2].sum()), self._bs, k)
                n_rec = be.choose(slot, norm, beta, 0, self.hi - self.lo)
                be.peer_zero_slab(slot)                   # ready for the next job once `check` has completed
                self.stats.append({"n_clusters": int(sc[:, k, 0].sum()), "max_len": int(sc[:, k, 1].max()),
                                   "n_records_local": int(n_rec), "updates": int(sc[:, k, 2].sum())})
        inner = self._inner
        inner.timer, inner.n_strands = self.timer, len(self._pending)
        out = inner.finish(v, m, tax_rank, higher, gather=gather)
        with self.timer("check"):
            parts = [self.torch.empty_like(self._bs) for _ in range(self.world)]
            self.dist.all_gather(parts, self._bs)
            bs = self.torch.stack(parts).cpu().numpy().reshape(self.world, -1)
            self._dirty = False
            for k in range(len(self._pending)):
                if int(bs[:, k].sum()) != int(sc[:, k, 3].sum()) and not os.environ.get("LIME_DEBUG"):
                    raise CarryDetected(f"strand {k}: byte sum {int(bs[:, k].sum())} != total added {int(sc[:, k, 3].sum())}")
        return out

    def close(self):
        """Collective: unmaps and frees the slabs; the backend then allocates whole matrices again."""
        if self.closed:
            return
        self.closed = True
        self.dist.barrier()
        self.be.peer_close()


def run_sharded(dist, torch, backend, rank, world, n_reads, n_refs, shards, alpha, beta, ebwt_mode, read_len,
                v, m=None, tax_rank=1, higher=False, gather=True, peer_job=None):
    """One classification job over `world` ranks: through `peer_job` (a PeerJob, reused between calls)
    when given, and again through the exchange route when that one reports a cell above 255 -- every rank
    sees the same all-gathered sums, so all ranks take the same branch; the peer job is closed then (its
    slabs are released so that the backend allocates whole matrices again).  shards: this rank's shard
    of every strand (dicts as for run_strand).  Returns (assignments, per-strand stats, route)."""
    if peer_job is not None and not peer_job.closed:
        try:
            peer_job.begin()
            for s, sh in enumerate(shards):
                peer_job.run_strand(s, sh, alpha, beta, ebwt_mode, read_len)
            return peer_job.finish(v, m, tax_rank, higher, gather=gather), list(peer_job.stats), "peer"
        except CarryDetected:
            peer_job.close()
    job = ShardedJob(dist, torch, backend, rank, world, n_reads, n_refs)
    for s, sh in enumerate(shards):
        job.run_strand(s, sh, alpha, beta, ebwt_mode, read_len)
    return job.finish(v, m, tax_rank, higher, gather=gather), list(job.stats), "exchange"


class _NoTimer:
    def __enter__(self):
        return self

    def __exit__(self, *a):
        return False
