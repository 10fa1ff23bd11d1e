"""CPU coverage of the multi-GPU host logic (lime_b200/multi.py): world size 2 (and 3) over gloo.
The per-rank compute is played by a checker backend built on the oracle (test infrastructure); what
is under test is the sharding plan, the halo exchange, the routing of (read, genome, count)
triples by read owner, the merge and the gather -- the result must equal the single-process
oracle result on the whole arrays, bit for bit."""
import os
import socket

import numpy as np
import pytest

import corpus as C


class OracleBackend:
    """Backend interface of lime_b200.multi on top of the oracle (numpy, CPU)."""
    device = None

    def __init__(self, oracle, torch):
        self.o, self.torch = oracle, torch
        self.slots = {}

    def load(self, lcp, da, ebwt, n_own, n_halo, pos_base, is_last, n_reads, n_refs):
        self.lcp, self.da = lcp.numpy().view(np.uint32), da.numpy().view(np.uint32)
        self.ebwt = None if ebwt is None else ebwt.numpy()
        self.n_own, self.n_tot, self.is_last = n_own, n_own + n_halo, is_last
        self.n_reads, self.n_refs = n_reads, n_refs
        self.sim = np.zeros((n_reads, n_refs), np.uint8)

    def cluster_similarity(self, alpha, ebwt_mode):
        lcp, da = self.lcp[:self.n_tot], self.da[:self.n_tot]
        cl, _ = self.o.cluster_detect(lcp, da, self.n_reads, alpha)
        cl = cl[cl["pStart"] < self.n_own]                      # ownership: head inside the shard
        if not self.is_last and cl.shape[0]:
            assert int(cl["pStart"][-1] + cl["len"][-1]) < self.n_tot, "halo too short for this test"
        self.sim = self.o.similarity(cl, da, self.ebwt[:self.n_tot] if ebwt_mode else None, self.n_reads, self.n_refs)
        return int(cl.shape[0]), int(cl["len"].max()) if cl.shape[0] else 0

    def updates(self):
        return int(np.count_nonzero(self.sim))

    def new_buffer(self, n):
        return self.torch.zeros(max(int(n), 1), dtype=self.torch.int64)

    def export(self, lo, hi, buf, offset):
        r, g = np.nonzero(self.sim[lo:hi])
        c = self.sim[lo:hi][r, g].astype(np.int64)
        packed = (r.astype(np.int64) + lo) | (((g.astype(np.int64) << 8) | c) << 32)
        buf[offset:offset + packed.shape[0]] = self.torch.from_numpy(packed)
        return int(packed.shape[0])

    def import_(self, buf, offset, n):
        t = buf[offset:offset + n].numpy()
        r, y = t & 0xFFFFFFFF, t >> 32
        keys = r * (1 << 32) + (y >> 8)
        assert np.unique(keys).shape[0] == n                    # unique within one source
        self.sim[r, y >> 8] = (self.sim[r, y >> 8].astype(np.int64) + (y & 0xFF)).astype(np.uint8)

    def choose(self, slot, norm, beta, lo, hi):
        self.slots[slot] = self.o.choose(self.sim[lo:hi], norm, np.float32(beta))
        return int(self.slots[slot][0].shape[0])

    def classify(self, n_files, n_refs, v, m, tax_rank, higher):
        bins = [self.slots[i][0] for i in range(n_files)]
        poss = [self.slots[i][1] for i in range(n_files)]
        return self.o.classify(bins, poss, n_refs, v, m, tax_rank, higher)


    # ---- emulation of the record exchange (XchgJob): every rank's inbox is one shared file, a source
    # writes its records {row, genome << 8 | count} into its own region of the owner's file (what
    # route_kernel does over NVLink), then its record count and a flag (the number of the accumulation) into the
    # owner's mailbox (publish_kernel); the owner polls the flags of all sources (wait_kernel) -- no collective
    def xchg_setup(self, dist, world, rank, per, n_refs, n_strands, flat_cap, inbox_cap):
        self.xd = (dist, world, rank, per, n_strands, inbox_cap)
        path = lambda r: os.path.join(self.shm_dir, f"inbox{r}.i64")
        mpath = lambda r: os.path.join(self.shm_dir, f"mail{r}.i64")
        np.memmap(path(rank), dtype=np.int64, mode="w+", shape=(n_strands, world, inbox_cap)).flush()
        np.memmap(mpath(rank), dtype=np.int64, mode="w+", shape=(n_strands, world, 2)).flush()
        dist.barrier()
        self.inbox = [np.memmap(path(r), dtype=np.int64, mode="r+", shape=(n_strands, world, inbox_cap)) for r in range(world)]
        self.mail = [np.memmap(mpath(r), dtype=np.int64, mode="r+", shape=(n_strands, world, 2)) for r in range(world)]
        self.x_state, self.epoch = {}, [0] * n_strands

    def xchg_accumulate(self, alpha, ebwt_mode, strand):
        dist, world, rank, per, n_strands, cap = self.xd
        n_cl, mx = self.cluster_similarity(alpha, ebwt_mode)
        over = False
        self.epoch[strand] += 1
        for k in range(world):
            lo, hi = k * per, min((k + 1) * per, self.n_reads)
            n = 0
            if hi > lo:
                r, g = np.nonzero(self.sim[lo:hi])
                c = self.sim[lo:hi][r, g].astype(np.int64)
                rec = r.astype(np.int64) | (((g.astype(np.int64) << 8) | c) << 32)
                n = rec.shape[0]
                over |= n > cap
                self.inbox[k][strand, rank, :min(n, cap)] = rec[:cap]
                self.inbox[k].flush()
            self.mail[k][strand, rank, 0] = n                   # the count before the flag
            self.mail[k].flush()
            self.mail[k][strand, rank, 1] = self.epoch[strand]
            self.mail[k].flush()
        self.x_state[strand] = dict(n_clusters=n_cl, max_len=mx, updates=int(np.count_nonzero(self.sim)), list_slots=0,
                                    err=4 if over else 0)

    def xchg_apply(self, strand):
        import time
        dist, world, rank, per, n_strands, cap = self.xd
        mine = np.memmap(self.inbox[rank].filename, dtype=np.int64, mode="r", shape=(n_strands, world, cap))
        slab = np.zeros((per, self.n_refs), np.int64)
        st = self.x_state[strand]
        for p in range(world):
            t0 = time.time()
            while True:
                mail = np.memmap(self.mail[rank].filename, dtype=np.int64, mode="r", shape=(n_strands, world, 2))
                if int(mail[strand, p, 1]) >= self.epoch[strand]:
                    break
                if time.time() - t0 > 20:
                    st["err"] |= 16
                    break
                time.sleep(0.001)
            t = np.array(mine[strand, p, :min(int(mail[strand, p, 0]), cap)])
            np.add.at(slab, (t & 0xFFFFFFFF, (t >> 40)), (t >> 32) & 0xFF)
        st["applied_t"] = int(slab.sum())
        self.sim = (slab & 0xFF).astype(np.uint8)               # the uint8 cell wraps
        st["byte_sum"] = int(self.sim.astype(np.int64).sum())

    def xchg_choose(self, strand, norm, beta, n_own):
        self.slots[strand] = self.o.choose(self.sim[:n_own], norm, np.float32(beta))
        self.x_state[strand]["n_records"] = int(self.slots[strand][0].shape[0])

    def xchg_status(self, strand):
        return dict(self.x_state[strand])

    def xchg_close(self):
        self.inbox, self.x_state = None, {}

    def classify_checked(self, n_files, n_refs, v, m, tax_rank, higher):
        """lime_classify after deferred strands: LIME_ERR_RETRY when a cell exceeded 255 or a list overflowed."""
        from lime_b200._lib import ERR_RETRY, LimeError
        for s, st in getattr(self, "x_state", {}).items():
            if st["applied_t"] != st["byte_sum"] or st["err"]:
                raise LimeError(ERR_RETRY, "deferred strand: a similarity cell exceeded 255")
        return self.classify_plain(n_files, n_refs, v, m, tax_rank, higher)


OracleBackend.classify_plain = OracleBackend.classify
OracleBackend.classify = OracleBackend.classify_checked


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, strands, n_reads, n_refs, read_len, alpha, beta, halo, v, m, q, peer_dir=None):
    import torch
    import torch.distributed as dist
    from oracle.oracle import Oracle
    from lime_b200 import multi
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    be = OracleBackend(Oracle(), torch)
    if peer_dir:
        be.shm_dir = peer_dir
        job = multi.XchgJob(dist, torch, be, rank, world, n_reads, n_refs, len(strands), 1 << 16, 1 << 15)
        job.begin()
    else:
        job = multi.ShardedJob(dist, torch, be, rank, world, n_reads, n_refs)
    job_shards = []
    for s, (ebwt, lcp, da) in enumerate(strands):
        n = lcp.shape[0]
        a, b = multi.shard_ranges(n, world)[rank]
        h = halo if rank < world - 1 else 0
        t = [torch.zeros(b - a + halo, dtype=torch.int32) for _ in range(2)] + [torch.zeros(b - a + halo, dtype=torch.uint8)]
        t[0][:b - a] = torch.from_numpy(lcp[a:b].view(np.int32))
        t[1][:b - a] = torch.from_numpy(da[a:b].view(np.int32))
        t[2][:b - a] = torch.from_numpy(ebwt[a:b])
        got = multi.exchange_halo(dist, rank, world, t, b - a, halo)
        assert got == h
        job_shards.append(dict(lcp=t[0], da=t[1], ebwt=t[2], n_own=b - a, n_halo=h, pos_base=a, is_last=rank == world - 1))
        job.run_strand(s, job_shards[-1], alpha, beta, 1, read_len)
    typ, tax, sim = job.finish(v, m, 1, m is not None)
    if peer_dir:
        # slabs and inboxes are reused by the next job
        first = [np.array(x) for x in (typ, tax, sim)]
        job.begin()
        for s, sh in enumerate(job_shards):
            job.run_strand(s, sh, alpha, beta, 1, read_len)
        typ, tax, sim = job.finish(v, m, 1, m is not None)
        assert all(a.tobytes() == np.asarray(b).tobytes() for a, b in zip(first, (typ, tax, sim)))
        # bench.py's schedule: a second context (backend) of this rank, job k + 1 queued before job k is finished
        be2 = OracleBackend(Oracle(), torch)
        be2.shm_dir = os.path.join(peer_dir, "second")
        os.makedirs(be2.shm_dir, exist_ok=True)
        job2 = multi.XchgJob(dist, torch, be2, rank, world, n_reads, n_refs, len(strands), 1 << 16, 1 << 15)
        for k_steps in (1, 2, 3):
            typ, tax, sim = multi.run_jobs_in_flight((job, job2), k_steps, job_shards, alpha, beta, 1, read_len, v, m, 1,
                                                     m is not None, gather=True)
            assert all(a.tobytes() == np.asarray(b).tobytes() for a, b in zip(first, (typ, tax, sim))), k_steps
        job2.close()
        job.close()
    q.put((rank, typ, tax, sim, job.stats))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("world,peer", [(2, False), (3, False), (2, True), (3, True)])
def test_sharded_job_equals_single_process(oracle, world, peer, tmp_path):
    import torch.multiprocessing as mp
    n_reads, n_refs, read_len, alpha, beta = 1201, 9, 100, 14, 0.25
    strands = []
    for s in range(2):
        c = C.synth_case(f"mg{s}", 180_001, n_reads, n_refs, seed=60 + s, p_long=1e-3, max_seg=600)
        strands.append((c["ebwt"], c["lcp"], c["da"]))
    v = (np.arange(n_refs, dtype=np.uint32) // 2) + 100
    m = np.tile(np.array([[1], [2], [3], [4], [5], [6]], np.uint32), (1, n_refs))
    # single-process oracle result
    bins, poss, scal = [], [], []
    for (ebwt, lcp, da) in strands:
        cl, mx = oracle.cluster_detect(lcp, da, n_reads, alpha)
        sim = oracle.similarity(cl, da, ebwt, n_reads, n_refs)
        b, p = oracle.choose(sim, read_len + 1 - alpha, np.float32(beta))
        bins.append(b); poss.append(p); scal.append((cl.shape[0], mx))
    exp = oracle.classify(bins, poss, n_refs, v, m, 1, True)
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, strands, n_reads, n_refs, read_len, alpha, beta, 2048, v, m, q,
                                               str(tmp_path) if peer else None))
             for r in range(world)]
    for p in procs:
        p.start()
    res = [q.get(timeout=300) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    for rank, typ, tax, sim, stats in res:
        assert typ.tobytes() == exp[0].tobytes() and tax.tobytes() == exp[1].tobytes() and sim.tobytes() == exp[2].tobytes()
        for s in range(2):
            assert (stats[s]["n_clusters"], stats[s]["max_len"]) == scal[s]
        assert peer or stats[0]["triples_sent"] > 0
    assert set(chr(t) for t in exp[0]) >= {"C", "U"}


def _worker_fallback(rank, world, port, case, alpha, beta, halo, v, q, peer_dir):
    import torch
    import torch.distributed as dist
    from oracle.oracle import Oracle
    from lime_b200 import multi
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    be = OracleBackend(Oracle(), torch)
    be.shm_dir = peer_dir
    n = case["lcp"].shape[0]
    a, b = multi.shard_ranges(n, world)[rank]
    t = [torch.zeros(b - a + halo, dtype=torch.int32) for _ in range(2)] + [torch.zeros(b - a + halo, dtype=torch.uint8)]
    t[0][:b - a] = torch.from_numpy(case["lcp"][a:b].view(np.int32))
    t[1][:b - a] = torch.from_numpy(case["da"][a:b].view(np.int32))
    t[2][:b - a] = torch.from_numpy(case["ebwt"][a:b])
    h = multi.exchange_halo(dist, rank, world, t, b - a, halo)
    shard = dict(lcp=t[0], da=t[1], ebwt=t[2], n_own=b - a, n_halo=h, pos_base=a, is_last=rank == world - 1)
    pj = multi.XchgJob(dist, torch, be, rank, world, case["n_reads"], case["n_refs"], 2, 1 << 12, 1 << 10)
    out, stats, route = multi.run_sharded(dist, torch, be, rank, world, case["n_reads"], case["n_refs"], [shard, shard],
                                          alpha, beta, 1, case["read_len"], v, xchg_job=pj)
    pj.close()
    q.put((rank, out, stats, route))
    dist.barrier()
    dist.destroy_process_group()


def test_xchg_job_falls_back_when_a_cell_exceeds_255(oracle, tmp_path):
    """Cell sums above 255 (the uint8 wrap of SimArray_): the byte-sum proof of the record-exchange job fails
    (the ranks agree on it by an all-reduce) and run_sharded redoes the job through the exact route, whose
    result wraps like the reference."""
    import torch.multiprocessing as mp
    case = C.wrap_case(n=30000)
    alpha, beta, world = 16, 0.0, 2
    v = np.arange(case["n_refs"], dtype=np.uint32) + 7
    cl, mx = oracle.cluster_detect(case["lcp"], case["da"], case["n_reads"], alpha)
    sim = oracle.similarity(cl, case["da"], case["ebwt"], case["n_reads"], case["n_refs"])
    b, p = oracle.choose(sim, case["read_len"] + 1 - alpha, np.float32(beta))
    exp = oracle.classify([b, b], [p, p], case["n_refs"], v, None, 1, False)
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker_fallback, args=(r, world, port, case, alpha, beta, int(mx) + 8, v, q, str(tmp_path)))
             for r in range(world)]
    for pr in procs:
        pr.start()
    res = [q.get(timeout=300) for _ in range(world)]
    for pr in procs:
        pr.join(timeout=60)
        assert pr.exitcode == 0
    for rank, out, stats, route in res:
        assert route == "exchange"
        assert all(np.asarray(g).tobytes() == e.tobytes() for g, e in zip(out, exp))
        assert (stats[0]["n_clusters"], stats[0]["max_len"]) == (cl.shape[0], mx)


def test_owner_division_by_multiply_high():
    """owner_of() of csrc/xchg.cu finds a read's owner with q = umulhi(r, magic), magic =
    min(floor(2^32 / per), 2^32 - 1), and one correction step; restated here and checked against
    integer division over the whole 32-bit range of read ids for awkward divisors."""
    rng = np.random.default_rng(11)
    pers = [1, 2, 3, 5, 7, 255, 256, 257, 65535, 65536, 65537, 1_000_003, 2_000_000, 2 ** 31 - 1, 2 ** 31, 2 ** 32 - 1]
    pers += [int(x) for x in rng.integers(1, 2 ** 32, size=40)]
    for per in pers:
        magic = min((1 << 32) // per, (1 << 32) - 1)
        r = np.concatenate([rng.integers(0, 1 << 32, size=20000, dtype=np.uint64),
                            np.arange(0, min(4 * per, 1 << 32), max(per // 3, 1), dtype=np.uint64)[:20000],
                            np.array([0, per - 1, per, per + 1 if per + 1 < (1 << 32) else 0, (1 << 32) - 1], dtype=np.uint64)])
        for edge in (per, 2 * per, 3 * per):
            if edge < (1 << 32):
                r = np.concatenate([r, np.array([edge - 1, edge], dtype=np.uint64)])
        q = (r * np.uint64(magic)) >> np.uint64(32)
        rem = (r - q * np.uint64(per)) & np.uint64(0xFFFFFFFF)            # 32-bit arithmetic of the kernel
        fix = rem >= np.uint64(per)
        q = q + fix.astype(np.uint64)
        rem = np.where(fix, rem - np.uint64(per), rem)
        assert np.array_equal(q, r // np.uint64(per)), per
        assert np.array_equal(rem, r % np.uint64(per)), per
