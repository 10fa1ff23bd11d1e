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
