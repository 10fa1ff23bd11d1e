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


def _check_case(ctx, oracle, case, alpha, ebwt_mode, beta=0.25):
    ebwt = case["ebwt"] if ebwt_mode else None
    ctx.load(case["lcp"], case["da"], ebwt, case["n_reads"], case["n_refs"])
    n_cl, max_len = ctx.cluster_detect(alpha)
    cl_o, max_o = oracle.cluster_detect(case["lcp"], case["da"], case["n_reads"], alpha)
    assert (n_cl, max_len) == (cl_o.shape[0], max_o), case["name"]
    assert np.array_equal(ctx.clusters(), cl_o), case["name"]
    if max_len > 65536:
        from lime_b200.api import LimeError
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
def test_steps_1_2_match_oracle(ctx, oracle, corpus, alpha, ebwt_mode):
    saw_exact = False
    for case in corpus:
        st = _check_case(ctx, oracle, case, alpha, ebwt_mode, beta=0.0 if case["name"] == "wrap" else 0.25)
        if st and st["exact_pass"]:
            saw_exact = True
    assert saw_exact          # the wrap case must have exercised the exact 32-bit pass


def test_medium_synthetic(ctx, oracle):
    case = C.synth_case("synth_medium", 6_000_000, 30_000, 40, seed=31, p_long=2e-5, max_seg=60000)
    for alpha in (12, 16):
        st = _check_case(ctx, oracle, case, alpha, 1)
        assert st["warp_clusters"] > 0
    st = _check_case(ctx, oracle, case, 12, 0)
    assert st["block_clusters"] > 0


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
