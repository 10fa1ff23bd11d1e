"""GPU parity: the CUDA path, called through the C ABI, against the oracle on the same inputs.
Bit-exact for clusters, similarity cells, .res.bin / .res.pos images and assignments."""
import numpy as np
import pytest

import corpus as C
from test_oracle_vs_reference import _random_rows

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ctx():
    from lime_b200.api import LimeContext
    c = LimeContext(0)
    yield c
    c.close()


def _check_case(ctx, oracle, case, alpha, ebwt_mode, beta=0.25, fused=False, keep=True):
    """fused=False: lime_cluster_detect + lime_similarity (cluster list in between);
    fused=True: lime_cluster_similarity (one pass, records kept only when keep)."""
    from lime_b200.api import LimeError
    ebwt = case["ebwt"] if ebwt_mode else None
    ctx.load(case["lcp"], case["da"], ebwt, case["n_reads"], case["n_refs"])
    cl_o, max_o = oracle.cluster_detect(case["lcp"], case["da"], case["n_reads"], alpha)
    if fused:
        if max_o > 65536:
            with pytest.raises(LimeError) as e:
                ctx.cluster_similarity(alpha, ebwt_mode, keep)
            assert e.value.code == -3
            return None
        n_cl, max_len = ctx.cluster_similarity(alpha, ebwt_mode, keep)
        assert (n_cl, max_len) == (cl_o.shape[0], max_o), case["name"]
        if keep:
            assert np.array_equal(ctx.clusters(), cl_o), case["name"]
    else:
        n_cl, max_len = ctx.cluster_detect(alpha)
        assert (n_cl, max_len) == (cl_o.shape[0], max_o), case["name"]
        assert np.array_equal(ctx.clusters(), cl_o), case["name"]
        if max_len > 65536:
            with pytest.raises(LimeError) as e:
                ctx.similarity(ebwt_mode)
            assert e.value.code == -3
            return None
        ctx.similarity(ebwt_mode)
    sim_o = oracle.similarity(cl_o, case["da"], ebwt, case["n_reads"], case["n_refs"])
    sim = ctx.sim_matrix()
    assert np.array_equal(sim, sim_o), (case["name"], alpha, ebwt_mode, ctx.sim_stats())
    norm = (case["read_len"] + 1 - alpha) & 0xFFFFFFFF
    n_rec = ctx.choose(0, norm, beta)
    bin_g, pos_g = ctx.choose_result(0)
    bin_o, pos_o = oracle.choose(sim_o, norm, np.float32(beta))
    assert n_rec == bin_o.shape[0]
    assert bin_g.tobytes() == bin_o.tobytes(), case["name"]
    assert pos_g.tobytes() == pos_o.tobytes(), case["name"]
    return ctx.sim_stats()


@pytest.mark.parametrize("alpha", [8, 16])
@pytest.mark.parametrize("ebwt_mode", [1, 0])
@pytest.mark.parametrize("fused", [False, True])
def test_steps_1_2_match_oracle(ctx, oracle, corpus, alpha, ebwt_mode, fused):
    saw_exact = False
    for case in corpus:
        st = _check_case(ctx, oracle, case, alpha, ebwt_mode, beta=0.0 if case["name"] == "wrap" else 0.25,
                         fused=fused, keep=(alpha == 16))
        if st and st["exact_pass"]:
            saw_exact = True
    assert saw_exact          # the wrap case must have exercised the exact 32-bit pass


@pytest.mark.parametrize("fused", [False, True])
def test_medium_synthetic(ctx, oracle, fused):
    case = C.synth_case("synth_medium", 6_000_000, 30_000, 40, seed=31, p_long=2e-5, max_seg=60000)
    for alpha in (12, 16):
        st = _check_case(ctx, oracle, case, alpha, 1, fused=fused)
        assert st["warp_clusters"] > 0
    st = _check_case(ctx, oracle, case, 12, 0, fused=fused, keep=False)
    assert st["block_clusters"] > 0


@pytest.mark.parametrize("ebwt_mode", [1, 0])
def test_fused_tile_boundaries(ctx, oracle, ebwt_mode):
    """Many clusters of 20..400 positions: they straddle the 4096-position tiles, exceed the
    32-lane window and the 128-position look-ahead of the fused kernel in every combination."""
    case = C.synth_case("synth_longish", 1_500_000, 4000, 9, seed=41, p_long=2e-2, max_seg=1424)
    for alpha in (12, 14):
        st = _check_case(ctx, oracle, case, alpha, ebwt_mode, fused=True)
        assert st["block_clusters"] > 100 and st["warp_clusters"] > 1000


def test_fused_adversarial_layouts(ctx, oracle):
    """Hand-made head patterns around tile (4096) / word (32) / look-ahead (128) boundaries."""
    rng = np.random.default_rng(5)
    n, n_reads, n_refs = 3 * 4096 + 77, 50, 4
    for trial in range(12):
        lcp = np.full(n, 20, np.uint32)
        heads = set([0])
        for b in (4096, 8192, 12288):
            for off in rng.integers(-140, 140, size=6):
                heads.add(int(min(max(b + off, 0), n - 1)))
        for h in rng.integers(0, n, size=[0, 40, 400, 3000][trial % 4]):
            heads.add(int(h))
        if trial % 3 == 0:
            heads |= set(range(4096 - 3, 4096 + 3)) | set(range(8192 + 120, 8192 + 135))
        lcp[sorted(heads)] = 3
        if trial == 11:
            lcp[0] = 20                       # leading run without a head (A2)
        da = rng.integers(0, n_reads + n_refs, size=n).astype(np.uint32)
        if trial % 2:
            da[rng.random(n) < 0.97] = 0      # mostly one read: mixed only where a genome shows up
        ebwt = np.frombuffer(bytes(rng.choice(list(b"ACGTN\0R$"), size=n).tolist()), dtype=np.uint8).copy()
        case = dict(name=f"adv{trial}", lcp=lcp, da=da, ebwt=ebwt, n_reads=n_reads, n_refs=n_refs, read_len=100)
        _check_case(ctx, oracle, case, 16, 1, beta=0.0, fused=True)
        _check_case(ctx, oracle, case, 16, 0, beta=0.0, fused=True, keep=False)


def test_stream_chunk_boundaries(ctx, oracle):
    """Head patterns around the stream kernel's chunk geometry: chunks start every 480 positions,
    hold 512 (one word of look-ahead); clusters of 28..40 positions straddle every boundary."""
    rng = np.random.default_rng(17)
    n, n_reads, n_refs = 40 * 480 + 123, 60, 5
    for trial in range(10):
        lcp = np.full(n, 20, np.uint32)
        heads = set([0])
        for c in range(1, 40):
            for off in rng.integers(-40, 40, size=3):
                heads.add(int(min(max(480 * c + off, 0), n - 1)))
            if trial % 2:
                heads |= {480 * c - 1, 480 * c, 480 * c + 31, 480 * c + 32}
        for h in rng.integers(0, n, size=[0, 100, 1500][trial % 3]):
            heads.add(int(h))
        lcp[sorted(heads)] = 3
        da = rng.integers(0, n_reads + n_refs, size=n).astype(np.uint32)
        if trial % 3 == 1:
            da[rng.random(n) < 0.95] = 1
        ebwt = np.frombuffer(bytes(rng.choice(list(b"ACGTN\0R$"), size=n).tolist()), dtype=np.uint8).copy()
        case = dict(name=f"chunk{trial}", lcp=lcp, da=da, ebwt=ebwt, n_reads=n_reads, n_refs=n_refs, read_len=100)
        _check_case(ctx, oracle, case, 16, 1, beta=0.0, fused=True)
        _check_case(ctx, oracle, case, 16, 0, beta=0.0, fused=True, keep=False)


@pytest.mark.parametrize("n_refs,ebwt_mode,alpha", [(40, 1, 16), (9, 0, 12), (100, 1, 16), (300, 0, 14)])
def test_export_import_roundtrip(ctx, oracle, n_refs, ebwt_mode, alpha):
    """lime_sim_export / lime_sim_import (the multi-GPU exchange): the triples of a row range are
    exactly the non-zero cells of the oracle's matrix, and importing them into a freshly loaded
    (zero) matrix restores those rows."""
    import torch
    case = C.synth_case(f"synth_x{n_refs}", 1_200_000, 12_000, n_refs, seed=51 + n_refs, p_long=2e-5, max_seg=60000)
    ebwt = case["ebwt"] if ebwt_mode else None
    ctx.load(case["lcp"], case["da"], ebwt, case["n_reads"], n_refs)
    ctx.cluster_similarity(alpha, ebwt_mode, False)
    cl, _ = oracle.cluster_detect(case["lcp"], case["da"], case["n_reads"], alpha)
    sim_o = oracle.similarity(cl, case["da"], ebwt, case["n_reads"], n_refs)
    assert np.array_equal(ctx.sim_matrix(), sim_o)
    lo, hi = 3000, 9500
    buf = torch.empty(int(np.count_nonzero(sim_o)) + 8, dtype=torch.int64, device="cuda:0")
    n = ctx.sim_export(lo, hi, buf.data_ptr(), buf.numel())
    tr = buf[:n].cpu().numpy().view(np.uint32).reshape(-1, 2)
    tr = tr[np.lexsort((tr[:, 1] >> 8, tr[:, 0]))]               # the export order is unspecified
    r, g, c = tr[:, 0], tr[:, 1] >> 8, tr[:, 1] & 0xFF
    rr, gg = np.nonzero(sim_o[lo:hi])
    assert n == rr.shape[0]
    assert np.array_equal(r, rr + lo) and np.array_equal(g, gg) and np.array_equal(c, sim_o[lo:hi][rr, gg])
    ctx.load(case["lcp"], case["da"], ebwt, case["n_reads"], n_refs)          # zero matrix
    ctx.sim_import(buf.data_ptr(), n)
    want = np.zeros_like(sim_o)
    want[lo:hi] = sim_o[lo:hi]
    assert np.array_equal(ctx.sim_matrix(), want)


def _pieces_per_warp(sim):
    """Largest number of non-zero 16-byte pieces among the 32-row groups of a matrix (the wide-row choose
    kernel stashes up to 256 per group and re-reads the rest from global memory)."""
    rows, refs = sim.shape
    pad = np.zeros((-rows % 32 + rows, -refs % 16 + refs), dtype=np.uint8)
    pad[:rows, :refs] = sim
    nz = pad.reshape(pad.shape[0], -1, 16).any(axis=2).sum(axis=1)
    return int(nz.reshape(-1, 32).sum(axis=1).max())


@pytest.mark.parametrize("n_refs", [129, 130, 512, 513, 930, 1000, 1024, 1025])
def test_choose_wide_rows_sparse_and_dense(ctx, oracle, n_refs):
    """lime_choose on matrices written cell by cell (lime_sim_import): 129..1024 genomes take the streaming
    kernel; a sparse band, a dense band (stash overflow -> re-read), empty rows, rows below beta, a ragged tail."""
    import torch
    rng = np.random.default_rng(n_refs)
    n_reads = 256 * 3 + 77
    # a load without clusters: zero matrix of the right shape, and lime_choose takes its single-pass kernel
    ctx.load(np.zeros(4000, dtype=np.uint32), rng.integers(0, n_reads + n_refs, size=4000).astype(np.uint32), None, n_reads, n_refs)
    assert ctx.cluster_similarity(16, 0, False) == (0, 0)
    sim = np.zeros((n_reads, n_refs), dtype=np.uint8)
    sparse = rng.random((n_reads, n_refs)) < 3.0 / n_refs
    sim[sparse] = rng.integers(1, 256, size=int(sparse.sum()), dtype=np.uint8)
    dense_rows = np.r_[40:75, 300:420, n_reads - 5:n_reads]
    d = rng.random((dense_rows.shape[0], n_refs)) < 0.6
    sim[dense_rows] = np.where(d, rng.integers(1, 256, size=d.shape, dtype=np.uint8), 0).astype(np.uint8)
    sim[100:140] = 0
    sim[200:230] = np.minimum(sim[200:230], 9)                                   # below beta
    sim[50, :] = 255
    sim[51, n_refs - 1] = 1
    assert _pieces_per_warp(sim) > 256
    r, g = np.nonzero(sim)
    tr = np.empty((r.shape[0], 2), dtype=np.uint32)
    tr[:, 0] = r
    tr[:, 1] = (g.astype(np.uint32) << 8) | sim[r, g]
    buf = torch.from_numpy(tr.view(np.int64).reshape(-1).copy()).to("cuda:0")
    ctx.sim_import(buf.data_ptr(), buf.numel())
    assert np.array_equal(ctx.sim_matrix(), sim)
    for norm, beta in ((85, 0.25), (135, 0.0), (85, 0.9)):
        n_rec = ctx.choose(0, norm, beta)
        bin_g, pos_g = ctx.choose_result(0)
        bin_o, pos_o = oracle.choose(sim, norm, np.float32(beta))
        assert n_rec == bin_o.shape[0]
        assert bin_g.tobytes() == bin_o.tobytes() and pos_g.tobytes() == pos_o.tobytes()
    assert np.array_equal(ctx.sim_matrix(), sim)                                  # lime_choose leaves the matrix alone


@pytest.mark.parametrize("ebwt_mode,alpha,world", [(1, 16, 4), (0, 12, 3), (1, 12, 1)])
def test_xchg_loopback(oracle, ebwt_mode, alpha, world):
    """lime_xchg_*: the fused kernel appending update records (K12 MODE_FLAT + block path), route_kernel sorting
    them by owner, apply_kernel adding them to the owner's slab -- here with `world` virtual owners on one GPU
    (loopback; real peers are checked under torchrun by test_gpu_multi.py).  Each virtual rank must end with
    exactly the oracle's rows of its reads, the carry proof (sum of t applied == byte sum), the records of
    clusterChoose and the assignments; slabs are reused by a second job without clearing."""
    import torch
    from lime_b200.api import LimeContext
    case = C.synth_case(f"synth_xchg{ebwt_mode}", 1_500_000, 15_000, 37, seed=91 + ebwt_mode, p_long=2e-5, max_seg=60000)
    n_reads, n_refs = case["n_reads"], case["n_refs"]
    ebwt = case["ebwt"] if ebwt_mode else None
    cl, mx = oracle.cluster_detect(case["lcp"], case["da"], n_reads, alpha)
    sim_o = oracle.similarity(cl, case["da"], ebwt, n_reads, n_refs)
    norm = (case["read_len"] + 1 - alpha) & 0xFFFFFFFF
    v = (np.arange(n_refs, dtype=np.uint32) // 3) + 50
    per = (n_reads + world - 1) // world
    for rank in range(world):
        lo, hi = min(rank * per, n_reads), min((rank + 1) * per, n_reads)
        bin_o, pos_o = oracle.choose(sim_o[lo:hi], norm, np.float32(0.25))
        exp = oracle.classify([bin_o, bin_o], [pos_o, pos_o], n_refs, v, None, 1, False)
        c = LimeContext(0)
        try:
            c.set_stream(torch.cuda.current_stream().cuda_stream)
            c.xchg_alloc(world, rank, per, n_refs, 2, 4_000_000, 300_000)
            c.xchg_open(None)
            for rep in range(2):
                for strand in (0, 1):
                    c.load(case["lcp"], case["da"], ebwt, n_reads, n_refs)
                    c.xchg_accumulate(alpha, ebwt_mode, strand)
                    c.xchg_apply(strand)
                    slab = c.sim_matrix(rows=per)
                    assert np.array_equal(slab[:hi - lo], sim_o[lo:hi]) and not slab[hi - lo:].any()
                    c.xchg_choose(strand, norm, 0.25, hi - lo)
                got = c.classify(2, n_refs, v, None, 1, False)
                for g, e in zip(got, exp):
                    assert np.asarray(g).tobytes() == e.tobytes()
                for strand in (0, 1):
                    st = c.xchg_status(strand)
                    assert (st["n_clusters"], st["max_len"]) == (cl.shape[0], mx) and st["err"] == 0
                    assert st["applied_t"] == st["byte_sum"] == int(sim_o[lo:hi].astype(np.int64).sum())
                    assert st["n_records"] == bin_o.shape[0] and st["list_slots"] <= 4_000_000
                    assert c.strand_status(strand)[:3] == (cl.shape[0], mx, bin_o.shape[0])
                    bin_g, pos_g = c.choose_result(strand)
                    assert bin_g.tobytes() == bin_o.tobytes() and pos_g.tobytes() == pos_o.tobytes()
                    assert st["updates"] >= np.count_nonzero(sim_o)
            c.xchg_close()
        finally:
            c.close()


def test_xchg_reports_cells_above_255_and_short_lists(oracle):
    """The wrap corpus case through the record exchange: the carry proof fails (sum of t applied != byte sum of
    the slab) and lime_classify answers LIME_ERR_RETRY; the same for an update list that is too small."""
    import torch
    from lime_b200.api import LimeContext, LimeError
    case = C.wrap_case()
    v = np.arange(case["n_refs"], dtype=np.uint32)
    for flat_cap, why in ((1 << 20, "255"), (256, "overflow")):
        c = LimeContext(0)
        try:
            c.set_stream(torch.cuda.current_stream().cuda_stream)
            c.xchg_alloc(1, 0, case["n_reads"], case["n_refs"], 2, flat_cap, 1 << 20)
            c.xchg_open(None)
            for strand in (0, 1):
                c.load(case["lcp"], case["da"], case["ebwt"], case["n_reads"], case["n_refs"])
                c.xchg_accumulate(16, 1, strand)
                c.xchg_apply(strand)
                c.xchg_choose(strand, 85, 0.0, case["n_reads"])
            with pytest.raises(LimeError) as e:
                c.classify(2, case["n_refs"], v, None, 1, False)
            assert e.value.code == -8 and why in str(e.value)
        finally:
            c.close()


def test_deferred_strands_equal_the_synchronous_calls(ctx, oracle):
    """lime_strand_run (no host synchronisation, one choose pass that also sums the bytes and clears the
    matrix) against lime_cluster_similarity + lime_choose, on the narrow- and the wide-row choose kernels;
    the wrap case must come back as LIME_ERR_RETRY from lime_classify and run_job must then redo it exactly."""
    from lime_b200.api import LimeError
    from lime_b200.pipeline import run_job
    for n_refs, ebwt_mode, alpha in ((40, 1, 16), (300, 0, 12), (130, 1, 16)):
        cases = [C.synth_case(f"synth_def{n_refs}_{s}", 900_000 + 77 * s, 9_000, n_refs, seed=310 + n_refs + s, p_long=2e-5, max_seg=30000)
                 for s in range(2)]
        v = (np.arange(n_refs, dtype=np.uint32) // 2) + 9
        strands = [{"ebwt": c["ebwt"], "lcp": c["lcp"], "da": c["da"], "n": c["lcp"].shape[0]} for c in cases]
        exp, scal = _expected_job(oracle, cases, 9_000, n_refs, 100, alpha, 0.25, ebwt_mode, v, None, 1, False)
        for rep in range(3):                      # the first strand of a context sizes the buffers, then deferred
            got, stats = run_job(ctx, strands, 9_000, n_refs, 100, alpha, 0.25, ebwt_mode, v, None, 1, False, deferred=True)
            assert [tuple(int(x) for x in st) for st in stats] == [tuple(int(x) for x in sc) for sc in scal]
            for g, e in zip(got, exp):
                assert np.asarray(g).tobytes() == e.tobytes()
            for s, c in enumerate(cases):
                cl, _ = oracle.cluster_detect(c["lcp"], c["da"], 9_000, alpha)
                b, p = oracle.choose(oracle.similarity(cl, c["da"], c["ebwt"] if ebwt_mode else None, 9_000, n_refs),
                                     (100 + 1 - alpha), np.float32(0.25))
                bg, pg = ctx.choose_result(s)
                assert bg.tobytes() == b.tobytes() and pg.tobytes() == p.tobytes()
    # cells above 255: the deferred strand is refused, the job falls back to the exact path
    w = C.wrap_case()
    vw = np.arange(w["n_refs"], dtype=np.uint32)
    ws = [{"ebwt": w["ebwt"], "lcp": w["lcp"], "da": w["da"], "n": w["lcp"].shape[0]}] * 2
    expw, _ = _expected_job(oracle, [w, w], w["n_reads"], w["n_refs"], 100, 16, 0.0, 1, vw, None, 1, False)
    ctx.load(w["lcp"], w["da"], w["ebwt"], w["n_reads"], w["n_refs"])
    ctx.strand_run(0, 16, 1, 85, 0.0)
    ctx.load(w["lcp"], w["da"], w["ebwt"], w["n_reads"], w["n_refs"])
    ctx.strand_run(1, 16, 1, 85, 0.0)
    with pytest.raises(LimeError) as e:
        ctx.classify(2, w["n_refs"], vw, None, 1, False)
    assert e.value.code == -8
    gotw, _ = run_job(ctx, ws, w["n_reads"], w["n_refs"], 100, 16, 0.0, 1, vw, None, 1, False, deferred=True)
    for g, e2 in zip(gotw, expw):
        assert np.asarray(g).tobytes() == e2.tobytes()


def test_two_jobs_in_flight_on_two_contexts(oracle):
    """bench.py's schedule: two contexts on two streams, the strands of job k + 1 are queued (lime_load of device
    arrays + lime_strand_run) before lime_classify of job k is called.  Every job must equal the oracle's."""
    import torch
    from lime_b200.api import LimeContext, MEM_DEVICE
    n_reads, n_refs, alpha, read_len = 9_000, 40, 16, 100
    cases = [C.synth_case(f"inflight_{s}", 700_000 + 31 * s, n_reads, n_refs, seed=520 + s, p_long=2e-5, max_seg=30000) for s in range(4)]
    v = (np.arange(n_refs, dtype=np.uint32) // 2) + 9
    jobs = [(cases[0], cases[1]), (cases[2], cases[3]), (cases[1], cases[2])]        # three different jobs
    exp = [_expected_job(oracle, list(j), n_reads, n_refs, read_len, alpha, 0.25, 1, v, None, 1, False)[0] for j in jobs]
    dev = {id(c): tuple(torch.from_numpy(c[k].view(np.int32) if c[k].dtype == np.uint32 else c[k]).to("cuda:0") for k in ("lcp", "da", "ebwt"))
           for c in cases}
    ctxs, streams = [LimeContext(0), LimeContext(0)], [torch.cuda.Stream(), torch.cuda.Stream()]
    try:
        for c, st in zip(ctxs, streams):
            c.set_stream(st.cuda_stream)
        torch.cuda.synchronize()

        def launch(c, job):
            for s, case in enumerate(job):
                l, d, e = dev[id(case)]
                c.load(l.data_ptr(), d.data_ptr(), e.data_ptr(), n_reads, n_refs, mem=MEM_DEVICE, n_total=case["lcp"].shape[0])
                c.strand_run(s, alpha, 1, read_len + 1 - alpha, 0.25)

        for rep in range(3):
            order = [jobs[(rep + k) % 3] for k in range(6)]
            want = [exp[(rep + k) % 3] for k in range(6)]
            prev = None
            for k, job in enumerate(order):
                c = ctxs[k % 2]
                launch(c, job)
                if prev is not None:
                    got = prev[0].classify(2, n_refs, v, None, 1, False)
                    for g, e in zip(got, prev[1]):
                        assert np.asarray(g).tobytes() == e.tobytes(), (rep, k)
                prev = (c, want[k])
            got = prev[0].classify(2, n_refs, v, None, 1, False)
            for g, e in zip(got, prev[1]):
                assert np.asarray(g).tobytes() == e.tobytes()
    finally:
        torch.cuda.synchronize()
        for c in ctxs:
            c.close()


def test_sharded_loads_union_equals_whole(ctx, oracle):
    """Position shards with a halo (lime_load n_own / n_halo / pos_base / is_last): the clusters of
    the shards are a partition of the whole array's clusters and the matrices add up mod 256."""
    from lime_b200.api import LimeError
    case = C.synth_case("synth_shard", 900_000, 3000, 6, seed=43, p_long=1e-3, max_seg=700)
    lcp, da, ebwt = case["lcp"], case["da"], case["ebwt"]
    n = lcp.shape[0]
    alpha = 14
    cl_o, max_o = oracle.cluster_detect(lcp, da, case["n_reads"], alpha)
    sim_o = oracle.similarity(cl_o, da, ebwt, case["n_reads"], case["n_refs"])
    cuts = [0, 300_001, 300_001 + 4096 * 50, n]
    halo = 4096
    cls, total, mx = [], np.zeros_like(sim_o, dtype=np.uint32), 0
    for k in range(3):
        a, b = cuts[k], cuts[k + 1]
        h = min(halo, n - b)
        sl = slice(a, b + h)
        ctx.load(lcp[sl], da[sl], ebwt[sl], case["n_reads"], case["n_refs"], n_own=b - a, n_halo=h,
                 pos_base=a, is_last=(k == 2))
        n_cl, max_len = ctx.cluster_similarity(alpha, 1, True)
        cls.append(ctx.clusters()); mx = max(mx, max_len)
        total += ctx.sim_matrix()
    got = np.concatenate(cls)
    assert np.array_equal(got, cl_o) and mx == max_o
    assert np.array_equal((total & 0xFF).astype(np.uint8), sim_o)
    # a halo that is too short for a straddling cluster is an error, not a silent truncation
    c = cl_o[int(np.argmax(cl_o["len"]))]
    assert c["len"] >= 12
    cut, h = int(c["pStart"]) + 3, int(c["len"] - 3) // 2
    ctx.load(lcp[:cut + h], da[:cut + h], ebwt[:cut + h], case["n_reads"], case["n_refs"], n_own=cut, n_halo=h,
             pos_base=0, is_last=False)
    with pytest.raises(LimeError) as e:
        ctx.cluster_similarity(alpha, 1, False)
    assert e.value.code == -4


def test_clusters_set_any_order(ctx, oracle, corpus):
    case = corpus[4]
    cl_o, _ = oracle.cluster_detect(case["lcp"], case["da"], case["n_reads"], 16)
    rng = np.random.default_rng(0)
    ctx.load(case["lcp"], case["da"], case["ebwt"], case["n_reads"], case["n_refs"])
    ctx.clusters_set(cl_o[rng.permutation(cl_o.shape[0])])
    ctx.similarity(1)
    assert np.array_equal(ctx.sim_matrix(), oracle.similarity(cl_o, case["da"], case["ebwt"], case["n_reads"], case["n_refs"]))


@pytest.mark.parametrize("n_files", [2, 4])
@pytest.mark.parametrize("higher", [False, True])
@pytest.mark.parametrize("tax_rank", [1, 2])
def test_step3_matches_oracle(ctx, oracle, tmp_path, n_files, higher, tax_rank):
    from lime_b200 import api
    rng = np.random.default_rng(7 + n_files + 10 * higher + tax_rank)
    n_reads, n_refs = 5000, 12
    lin = str(tmp_path / "lineage.csv")
    with open(lin, "w", newline="") as f:
        f.write(C.lineage_text(C.default_lineage(n_refs, holes=True, seed=3), crlf=True, trailing_blank=True))
    v, m, found = api.fix_rank(lin, tax_rank, n_refs, higher)
    vo, mo, fo = oracle.fix_rank(lin, tax_rank, n_refs, higher)
    assert np.array_equal(v, vo) and found == fo and (m is None or np.array_equal(m, mo))
    bins, poss = [], []
    for i in range(n_files):
        b, p = _random_rows(rng, n_reads, n_refs)
        ctx.slot_set(i, b, p)
        bins.append(b); poss.append(p)
    typ, tax, sim = ctx.classify(n_files, n_refs, v, m, tax_rank, higher)
    typ_o, tax_o, sim_o = oracle.classify(bins, poss, n_refs, vo, mo, tax_rank, higher)
    assert np.array_equal(typ, typ_o)
    assert np.array_equal(tax, tax_o)
    assert sim.tobytes() == sim_o.tobytes()
    a, b = str(tmp_path / "gpu.csv"), str(tmp_path / "oracle.csv")
    api.write_csv(a, typ, tax, sim)
    oracle.write_csv(b, typ_o, tax_o, sim_o)
    assert open(a, "rb").read() == open(b, "rb").read()


def test_step3_tiny_sims_all_genomes_corner(ctx, oracle):
    """h < ERROR: every genome of the database is 'near the maximum' in Exam_2 (Classify.cpp:230-253)."""
    n_refs = 5
    recs = np.zeros(5, dtype=[("sim", "<f4"), ("idRef", "<u4")])
    recs[1] = (0.01, 2); recs[2] = (0.01, 1); recs[3] = (0.005, 3); recs[4] = (0.0, 0)
    pos = np.array([1, 0], dtype=np.uint64)
    for v, m in ((np.array([7, 7, 8, 7, 7], np.uint32), None), (np.array([7, 8, 8, 9, 7], np.uint32), np.tile(np.array([[1], [2], [3], [4], [5], [6]], np.uint32), (1, 5)))):
        for nf in (2, 4):
            for s in range(nf):
                ctx.slot_set(s, recs, pos)
            got = ctx.classify(nf, n_refs, v, m, 1, m is not None)
            exp = oracle.classify([recs] * nf, [pos] * nf, n_refs, v, m, 1, m is not None)
            for g, e in zip(got, exp):
                assert g.tobytes() == e.tobytes()


def test_full_paired_pipeline_exact(ctx, oracle, tmp_path):
    """4 strands (mate1 F/RC, mate2 F/RC as in LiME_paired.sh) -> Classify 4, on exact suffix-sorted data."""
    from lime_b200 import api, datagen
    rng = np.random.default_rng(77)
    genomes = [datagen.random_dna(rng, 700)]
    g2 = list(genomes[0])
    for j in range(700):
        if rng.random() < 0.01:
            g2[j] = "ACGT"[int(rng.integers(4))]
    genomes += ["".join(g2), datagen.random_dna(rng, 650)]
    n_reads, read_len, alpha, beta = 80, 50, 16, 0.25
    m1, _ = datagen.sample_reads(rng, genomes, n_reads, read_len, err=0.02)
    m2, _ = datagen.sample_reads(rng, genomes, n_reads, read_len, err=0.02)
    strands = [m1, [datagen.revcomp(r) for r in m1], m2, [datagen.revcomp(r) for r in m2]]
    v = np.array([10, 10, 20], np.uint32)
    m = np.array([[10, 10, 20], [1, 1, 2], [5, 5, 5], [0, 0, 0], [6, 6, 6], [7, 7, 7]], np.uint32)
    bins, poss = [], []
    for s, reads in enumerate(strands):
        ebwt, lcp, da = datagen.exact_collection(reads, genomes)
        ctx.load(lcp, da, ebwt, n_reads, 3)
        ctx.cluster_detect(alpha)
        ctx.similarity(1)
        ctx.choose(s, read_len + 1 - alpha, beta)
        cl, _ = oracle.cluster_detect(lcp, da, n_reads, alpha)
        b, p = oracle.choose(oracle.similarity(cl, da, ebwt, n_reads, 3), read_len + 1 - alpha, np.float32(beta))
        bg, pg = ctx.choose_result(s)
        assert bg.tobytes() == b.tobytes() and pg.tobytes() == p.tobytes()
        bins.append(b); poss.append(p)
    for higher in (False, True):
        got = ctx.classify(4, 3, v, m if higher else None, 1, higher)
        exp = oracle.classify(bins, poss, 3, v, m if higher else None, 1, higher)
        for g, e in zip(got, exp):
            assert g.tobytes() == e.tobytes()
    assert set(chr(t) for t in exp[0]) >= {"C"}


def _expected_job(oracle, cases, n_reads, n_refs, read_len, alpha, beta, ebwt_mode, v, m, tax_rank, higher):
    bins, poss, scal = [], [], []
    for c in cases:
        cl, mx = oracle.cluster_detect(c["lcp"], c["da"], n_reads, alpha)
        sim = oracle.similarity(cl, c["da"], c["ebwt"] if ebwt_mode else None, n_reads, n_refs)
        b, p = oracle.choose(sim, (read_len + 1 - alpha) & 0xFFFFFFFF, np.float32(beta))
        bins.append(b); poss.append(p); scal.append((cl.shape[0], mx, b.shape[0]))
    return oracle.classify(bins, poss, n_refs, v, m, tax_rank, higher), scal


@pytest.mark.parametrize("shape", ["cfg3_ebwt0_4strands", "cfg4_higher_len150_trlcp32", "cfg5_alpha12", "cfg5_alpha20", "cfg5_alpha24",
                                   "cfg3_930genomes", "cfg4_1000genomes"])
def test_baseline_config_shapes(ctx, oracle, shape):
    """Reduced-size versions of BASELINE.json configs 3-5 (config 2 is the bench workload, config 1 the exact
    paired-end test above): whole jobs through the fused path, assignments and per-strand scalars bit-exact."""
    from lime_b200.pipeline import run_job, synthetic_lineage
    if shape == "cfg3_930genomes":
        # the row width of config 3 (930 genomes, lineage Datasets/Reference_database.csv has 930 rows): the wide-row
        # choose kernel, export / classify over 930-wide rows (ClusterBWT_DA.cpp:385-402, Classify.cpp:171-260)
        n, n_reads, n_refs, read_len, alpha, ebwt_mode, higher, n_str, kw = 1_200_000, 6_000, 930, 100, 16, 0, False, 4, {}
    elif shape == "cfg4_1000genomes":
        n, n_reads, n_refs, read_len, alpha, ebwt_mode, higher, n_str, kw = 1_200_000, 6_000, 1000, 150, 16, 1, True, 4, {"trlcp": 32}
    elif shape == "cfg3_ebwt0_4strands":
        n, n_reads, n_refs, read_len, alpha, ebwt_mode, higher, n_str, kw = 600_000, 20_000, 93, 100, 16, 0, False, 4, {}
    elif shape == "cfg4_higher_len150_trlcp32":
        n, n_reads, n_refs, read_len, alpha, ebwt_mode, higher, n_str, kw = 700_000, 12_000, 100, 150, 16, 1, True, 4, {"trlcp": 32}
    else:
        n, n_reads, n_refs, read_len, alpha, ebwt_mode, higher, n_str, kw = 500_000, 15_000, 50, 100, int(shape[10:]), 1, False, 2, {}
    rows, v, m = synthetic_lineage(n_refs)
    cases = [C.synth_case(f"{shape}_{s}", n + 13 * s, n_reads, n_refs, read_len=read_len, seed=700 + s, p_long=2e-5, max_seg=20000, **kw)
             for s in range(n_str)]
    exp, scal = _expected_job(oracle, cases, n_reads, n_refs, read_len, alpha, 0.25, ebwt_mode, v, m if higher else None, 1, higher)
    strands = [{"ebwt": c["ebwt"], "lcp": c["lcp"], "da": c["da"], "n": c["lcp"].shape[0]} for c in cases]
    got, stats = run_job(ctx, strands, n_reads, n_refs, read_len, alpha, 0.25, ebwt_mode, v, m if higher else None, 1, higher)
    assert [tuple(int(x) for x in st) for st in stats] == [tuple(int(x) for x in sc) for sc in scal]
    for g, e in zip(got, exp):
        assert np.asarray(g).tobytes() == e.tobytes()
    if alpha <= 16:
        assert int((exp[0] == ord("C")).sum()) > 0
