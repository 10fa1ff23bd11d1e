"""GPU: lime_set_alpha_limit / lime_set_alpha + lime_load(LIME_MEM_HOST) (lcp crosses PCIe as min(lcp, 255) bytes and stays that
way on the device: the stream kernel's LCP8 variants read the bytes, lime_b200/csrc/lime_abi_load.cu) gives the
same clusters, similarity matrix and records as the plain load and as the oracle."""
import numpy as np
import pytest

import corpus as C

pytestmark = [pytest.mark.gpu, pytest.mark.timeout(600)]

CHUNK = 8 << 20      # LIME_STAGE_CHUNK (common.cuh)


@pytest.fixture(scope="module")
def ctx():
    from lime_b200.api import LimeContext
    c = LimeContext(0)
    yield c
    c.close()


def _results(ctx, case, alpha, ebwt_mode, fused):
    ebwt = case["ebwt"] if ebwt_mode else None
    ctx.load(case["lcp"], case["da"], ebwt, case["n_reads"], case["n_refs"])
    if fused:
        counts = ctx.cluster_similarity(alpha, ebwt_mode, True)
    else:
        counts = ctx.cluster_detect(alpha)
        ctx.similarity(ebwt_mode)
    cl = ctx.clusters().copy()
    sim = ctx.sim_matrix().copy()
    n_rec = ctx.choose(0, (case["read_len"] + 1 - alpha) & 0xFFFFFFFF, 0.25)
    bin_g, pos_g = ctx.choose_result(0)
    return counts, cl, sim, n_rec, bin_g.tobytes(), pos_g.tobytes()


def _big_lcp(case, alpha, seed):
    """Values the byte cannot hold, on positions that are not heads (lcp >= alpha stays >= alpha)."""
    rng = np.random.default_rng(seed)
    lcp = case["lcp"].copy()
    if alpha > 200:
        # the synthetic lcp stays below 255: lift everything from 16 up so that 16 lands exactly on 255 (not a
        # head at alpha = 255) and put some heads on 254 (still heads)
        heads = np.flatnonzero(lcp < 16)
        lcp = np.where(lcp >= 16, lcp + 239, lcp).astype(np.uint32)
        lcp[heads[::5]] = 254
    inner = np.flatnonzero(lcp >= alpha)
    pick = inner[rng.integers(0, inner.size, inner.size // 7)]
    lcp[pick] = rng.integers(255, 2**32, pick.size, dtype=np.uint64).astype(np.uint32)
    lcp[inner[:4]] = np.array([255, 256, 0xFFFFFFFF, 0x80000000], dtype=np.uint32)
    out = dict(case)
    out["lcp"] = lcp
    return out


@pytest.mark.parametrize("n", [70_001, 2 * CHUNK + 12_345])
@pytest.mark.parametrize("ebwt_mode,fused,alpha", [(1, True, 16), (0, False, 12), (1, True, 255)])
def test_narrowed_load_equals_plain_load(ctx, oracle, n, ebwt_mode, fused, alpha):
    case = _big_lcp(C.synth_case("narrow", n, max(1000, n // 300), 40, seed=17 + ebwt_mode), alpha, 3)
    ctx.set_alpha_limit(0)
    plain = _results(ctx, case, alpha, ebwt_mode, fused)
    ctx.set_alpha_limit(alpha)
    try:
        nar = _results(ctx, case, alpha, ebwt_mode, fused)
        nar2 = _results(ctx, case, alpha, ebwt_mode, fused)          # stage buffers reused
    finally:
        ctx.set_alpha_limit(0)
    cl_o, max_o = oracle.cluster_detect(case["lcp"], case["da"], case["n_reads"], alpha)
    assert tuple(plain[0]) == (cl_o.shape[0], max_o)
    for got in (nar, nar2):
        assert tuple(got[0]) == tuple(plain[0])
        assert np.array_equal(got[1], cl_o)
        assert np.array_equal(got[2], plain[2])
        assert got[3:] == plain[3:]
    if n < 1_000_000:
        ebwt = case["ebwt"] if ebwt_mode else None
        assert np.array_equal(nar[2], oracle.similarity(cl_o, case["da"], ebwt, case["n_reads"], case["n_refs"]))


def test_alpha_above_255_is_refused_on_a_narrowed_strand(ctx):
    from lime_b200.api import LimeError
    case = C.synth_case("narrow_refuse", 70_001, 1000, 40, seed=23)
    ctx.set_alpha_limit(16)
    try:
        ctx.load(case["lcp"], case["da"], case["ebwt"], case["n_reads"], case["n_refs"])
        with pytest.raises(LimeError) as e:
            ctx.cluster_detect(256)
        assert e.value.code == -5
        with pytest.raises(LimeError) as e:
            ctx.cluster_similarity(300, 1, False)
        assert e.value.code == -5
        ctx.cluster_detect(255)                      # the limit itself is fine
    finally:
        ctx.set_alpha_limit(0)
    # a limit above 255 is a promise the byte cannot keep: the load stays plain and any alpha works
    ctx.set_alpha_limit(1000)
    try:
        ctx.load(case["lcp"], case["da"], case["ebwt"], case["n_reads"], case["n_refs"])
        ctx.cluster_detect(256)
    finally:
        ctx.set_alpha_limit(0)


@pytest.mark.parametrize("n", [70_001, 2 * CHUNK + 12_345])
@pytest.mark.parametrize("ebwt_mode,fused,alpha", [(1, True, 16), (0, False, 12), (1, True, 255), (0, True, 1)])
def test_bit_load_equals_plain_load(ctx, oracle, n, ebwt_mode, fused, alpha):
    """lime_set_alpha: lcp crosses as the bit lcp < alpha and is expanded to the bytes 0 / 255 on the device."""
    case = _big_lcp(C.synth_case("bits", n, max(1000, n // 300), 40, seed=19 + ebwt_mode), max(alpha, 2), 5)
    ctx.set_alpha(0)
    plain = _results(ctx, case, alpha, ebwt_mode, fused)
    ctx.set_alpha(alpha)
    try:
        got1 = _results(ctx, case, alpha, ebwt_mode, fused)
        got2 = _results(ctx, case, alpha, ebwt_mode, fused)         # stage buffers reused
    finally:
        ctx.set_alpha(0)
    cl_o, max_o = oracle.cluster_detect(case["lcp"], case["da"], case["n_reads"], alpha)
    assert tuple(plain[0]) == (cl_o.shape[0], max_o)
    for got in (got1, got2):
        assert tuple(got[0]) == tuple(plain[0])
        assert np.array_equal(got[1], cl_o)
        assert np.array_equal(got[2], plain[2])
        assert got[3:] == plain[3:]


def test_another_alpha_is_refused_on_a_bit_loaded_strand(ctx):
    from lime_b200.api import LimeError
    case = C.synth_case("bits_refuse", 70_001, 1000, 40, seed=29)
    ctx.set_alpha(16)
    try:
        ctx.load(case["lcp"], case["da"], case["ebwt"], case["n_reads"], case["n_refs"])
        for other in (15, 17, 300):
            with pytest.raises(LimeError) as e:
                ctx.cluster_detect(other)
            assert e.value.code == -5
        with pytest.raises(LimeError) as e:
            ctx.cluster_similarity(12, 1, False)
        assert e.value.code == -5
        ctx.cluster_detect(16)
    finally:
        ctx.set_alpha(0)
    # alpha above 255: the load stays plain and any alpha works
    ctx.set_alpha(1000)
    try:
        ctx.load(case["lcp"], case["da"], case["ebwt"], case["n_reads"], case["n_refs"])
        ctx.cluster_detect(256)
    finally:
        ctx.set_alpha(0)


def test_bit_load_keeps_da_whole_when_ids_need_more_than_three_bytes(ctx):
    """n_reads + n_refs > 2^24: lcp still crosses as bits, da as it is (ids up to 2^24 + 4 in the array)."""
    n, n_reads, n_refs = 200_003, (1 << 24) - 2, 7
    rng = np.random.default_rng(31)
    case = C.synth_case("bits_wide_ids", n, 5000, n_refs, seed=31)
    da = case["da"].copy()
    genome = da >= 5000
    da[genome] = da[genome] - 5000 + n_reads                       # genomes sit above 2^24 - 2
    reads = np.flatnonzero(~genome)
    da[reads[::3]] = rng.integers(1 << 23, n_reads, reads[::3].size).astype(np.uint32)
    wide = dict(case, da=da, n_reads=n_reads)
    ctx.set_alpha(0)
    plain = _results(ctx, wide, 16, 1, True)
    ctx.set_alpha(16)
    try:
        got = _results(ctx, wide, 16, 1, True)
    finally:
        ctx.set_alpha(0)
    assert tuple(got[0]) == tuple(plain[0]) and np.array_equal(got[1], plain[1])
    assert np.array_equal(got[2], plain[2]) and got[3:] == plain[3:]
    assert plain[0][0] > 0 and plain[2].any()                        # clusters were found and scored
