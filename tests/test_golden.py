"""Golden vectors: outputs of the UNMODIFIED reference executables (tests/golden/*.npz, made by
tests/golden/make_golden.py).  CPU: the oracle reproduces them.  GPU: the CUDA path, through the
C ABI, reproduces them -- parity with the reference itself, without the oracle in between."""
import glob
import os

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
STEPS12 = sorted(glob.glob(os.path.join(HERE, "golden", "steps12_*.npz")))
CLASSIFY = sorted(glob.glob(os.path.join(HERE, "golden", "classify_*.npz")))
CL_DT = np.dtype([("pStart", "<u8"), ("len", "<u8")])
PS_DT = np.dtype([("sim", "<f4"), ("idRef", "<u4")])


def test_fixtures_are_present():
    assert len(STEPS12) >= 10 and len(CLASSIFY) == 2


def _norm(g):
    return (int(g["read_len"]) + 1 - int(g["alpha"])) & 0xFFFFFFFF


@pytest.mark.parametrize("path", STEPS12, ids=lambda p: os.path.basename(p)[8:-4])
def test_oracle_reproduces_reference_steps12(oracle, path):
    g = np.load(path)
    cl, mx = oracle.cluster_detect(g["lcp"], g["da"], int(g["n_reads"]), int(g["alpha"]))
    assert cl.tobytes() == g["clrs"].tobytes()
    out = g["out"].tobytes()
    assert int.from_bytes(out[12:20], "little") == mx and int.from_bytes(out[20:28], "little") == cl.shape[0]
    for ebwt in (1, 0):
        sim = oracle.similarity(cl, g["da"], g["ebwt"] if ebwt else None, int(g["n_reads"]), int(g["n_refs"]))
        b, p = oracle.choose(sim, _norm(g), np.float32(g["beta"]))
        assert b.tobytes() == g[f"res_bin_{ebwt}"].tobytes() and p.tobytes() == g[f"res_pos_{ebwt}"].tobytes()


@pytest.mark.parametrize("path", CLASSIFY, ids=lambda p: os.path.basename(p)[:-4])
def test_oracle_reproduces_reference_classify(oracle, tmp_path, path):
    g = np.load(path)
    nf = 2 if "classify_2" in path else 4
    lin = str(tmp_path / "lineage.csv")
    open(lin, "wb").write(g["lineage"].tobytes())
    bins = [np.frombuffer(g[f"bin{i}"].tobytes(), PS_DT) for i in range(nf)]
    poss = [g[f"pos{i}"] for i in range(nf)]
    for higher in (0, 1):
        for tax_rank in (1, 2):
            v, m, _ = oracle.fix_rank(lin, tax_rank, int(g["n_refs"]), bool(higher))
            typ, tax, sim = oracle.classify(bins, poss, int(g["n_refs"]), v, m, tax_rank, bool(higher))
            o = str(tmp_path / "o.csv")
            oracle.write_csv(o, typ, tax, sim)
            assert open(o, "rb").read() == g[f"csv_h{higher}_t{tax_rank}"].tobytes()


@pytest.fixture(scope="module")
def ctx():
    from lime_b200.api import LimeContext
    c = LimeContext(0)
    yield c
    c.close()


@pytest.mark.gpu
@pytest.mark.parametrize("fused", [True, False])
@pytest.mark.parametrize("path", STEPS12, ids=lambda p: os.path.basename(p)[8:-4])
def test_cuda_reproduces_reference_steps12(ctx, path, fused):
    g = np.load(path)
    n_reads, n_refs, alpha = int(g["n_reads"]), int(g["n_refs"]), int(g["alpha"])
    out = g["out"].tobytes()
    for ebwt in (1, 0):
        ctx.load(g["lcp"], g["da"], g["ebwt"] if ebwt else None, n_reads, n_refs)
        if fused:
            n_cl, mx = ctx.cluster_similarity(alpha, ebwt, True)
        else:
            n_cl, mx = ctx.cluster_detect(alpha)
            ctx.similarity(ebwt)
        assert ctx.clusters().tobytes() == g["clrs"].tobytes()
        assert int.from_bytes(out[12:20], "little") == mx and int.from_bytes(out[20:28], "little") == n_cl
        ctx.choose(0, _norm(g), float(g["beta"]))
        b, p = ctx.choose_result(0)
        assert b.tobytes() == g[f"res_bin_{ebwt}"].tobytes() and p.tobytes() == g[f"res_pos_{ebwt}"].tobytes()


@pytest.mark.gpu
@pytest.mark.parametrize("path", CLASSIFY, ids=lambda p: os.path.basename(p)[:-4])
def test_cuda_reproduces_reference_classify(ctx, tmp_path, path):
    from lime_b200 import api
    g = np.load(path)
    nf = 2 if "classify_2" in path else 4
    lin = str(tmp_path / "lineage.csv")
    open(lin, "wb").write(g["lineage"].tobytes())
    for i in range(nf):
        ctx.slot_set(i, np.frombuffer(g[f"bin{i}"].tobytes(), PS_DT), g[f"pos{i}"])
    for higher in (0, 1):
        for tax_rank in (1, 2):
            v, m, _ = api.fix_rank(lin, tax_rank, int(g["n_refs"]), bool(higher))
            typ, tax, sim = ctx.classify(nf, int(g["n_refs"]), v, m, tax_rank, bool(higher))
            o = str(tmp_path / "o.csv")
            api.write_csv(o, typ, tax, sim)
            assert open(o, "rb").read() == g[f"csv_h{higher}_t{tax_rank}"].tobytes()


# ---- BASELINE config 1 stand-in: the real example reads against three pseudo-genomes, 4 strands, the reference's
# own outputs (tests/golden/make_config1.py, LiME_paired.sh defaults alpha=16 beta=0.25 TaxRank=1) -------------------
CONFIG1 = os.path.join(HERE, "golden", "config1_example.npz")


def _config1_known_answers(g, csv_bytes):
    """Every read of an accession listed in the lineage file is classified to that accession's species, the reads
    of the fourth accession (not in the database) stay unclassified."""
    lineage = g["lineage"].tobytes().decode()
    tax = {l.split(";")[0]: l.split(";")[1] for l in lineage.splitlines()[1:] if l.strip()}
    rows = [l.split(",") for l in csv_bytes.decode().splitlines()[1:]]
    assert len(rows) == int(g["n_reads"])
    for row, acc in zip(rows, g["acc_of"]):
        if str(acc) in tax:
            assert row[0] == "C" and row[2] == tax[str(acc)]
        else:
            assert row[0] == "U"


def test_oracle_reproduces_reference_on_the_example_reads(oracle, tmp_path):
    g = np.load(CONFIG1)
    n_reads, n_refs = int(g["n_reads"]), int(g["n_refs"])
    lin = str(tmp_path / "LineageFile.csv")
    open(lin, "wb").write(g["lineage"].tobytes())
    bins, poss = [], []
    for s in range(4):
        cl, mx = oracle.cluster_detect(g[f"lcp_{s}"], g[f"da_{s}"], n_reads, 16)
        assert cl.tobytes() == g[f"clrs_{s}"].tobytes()
        out = g[f"out_{s}"].tobytes()
        assert int.from_bytes(out[12:20], "little") == mx and int.from_bytes(out[20:28], "little") == cl.shape[0]
        b, p = oracle.choose(oracle.similarity(cl, g[f"da_{s}"], g[f"ebwt_{s}"], n_reads, n_refs), 100 + 1 - 16, np.float32(0.25))
        assert b.tobytes() == g[f"res_bin_{s}"].tobytes() and p.tobytes() == g[f"res_pos_{s}"].tobytes()
        bins.append(b); poss.append(p)
    v, m, found = oracle.fix_rank(lin, 1, n_refs, False)
    assert found == n_refs
    o = str(tmp_path / "o.csv")
    oracle.write_csv(o, *oracle.classify(bins, poss, n_refs, v, None, 1, False))
    assert open(o, "rb").read() == g["csv"].tobytes()
    _config1_known_answers(g, g["csv"].tobytes())


@pytest.mark.gpu
@pytest.mark.parametrize("deferred", [False, True])
def test_cuda_reproduces_reference_on_the_example_reads(ctx, tmp_path, deferred):
    from lime_b200 import api
    from lime_b200.pipeline import run_job
    g = np.load(CONFIG1)
    n_reads, n_refs = int(g["n_reads"]), int(g["n_refs"])
    lin = str(tmp_path / "LineageFile.csv")
    open(lin, "wb").write(g["lineage"].tobytes())
    v, m, found = api.fix_rank(lin, 1, n_refs, False)
    assert found == n_refs
    strands = [{"ebwt": g[f"ebwt_{s}"], "lcp": g[f"lcp_{s}"], "da": g[f"da_{s}"], "n": g[f"lcp_{s}"].shape[0]} for s in range(4)]
    for rep in range(2 if deferred else 1):
        out, stats = run_job(ctx, strands, n_reads, n_refs, 100, 16, 0.25, 1, v, None, 1, False, deferred=deferred)
        for s in range(4):
            ref_out = g[f"out_{s}"].tobytes()
            assert (int(stats[s][1]), int(stats[s][0])) == (int.from_bytes(ref_out[12:20], "little"), int.from_bytes(ref_out[20:28], "little"))
            b, p = ctx.choose_result(s)
            assert b.tobytes() == g[f"res_bin_{s}"].tobytes() and p.tobytes() == g[f"res_pos_{s}"].tobytes()
        o = str(tmp_path / "o.csv")
        api.write_csv(o, *out)
        assert open(o, "rb").read() == g["csv"].tobytes()
