"""The formats either side of the path, host code only -- no GPU needed:
  * <fasta>.<numSeq>.gesa -> .ebwt / .lcp / .da   (EGSAtoBCR.cpp), against the reference converter;
  * BIN=0 text results <fasta>.res.txt: our writer against the reference's ClusterBWT_DA built with
    -DBIN=0, our reader + the oracle's classification against the reference's Classify built with -DBIN=0.
The reference binaries come from oracle/_ref (oracle/Makefile)."""
import os
import subprocess

import numpy as np
import pytest

import corpus as C
from lime_b200 import api, datagen
from test_oracle_vs_reference import _random_rows

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BIN = os.path.join(ROOT, "lime_b200", "host", "bin")


def _need(reference, exe):
    if not os.path.exists(os.path.join(reference.dir, exe)):
        pytest.skip(f"oracle/_ref/{exe} not built")


def _gesa_bytes(ebwt, lcp, da, rng):
    rec = np.zeros(ebwt.shape[0], dtype=np.dtype([("text", "<u4"), ("suff", "<u4"), ("lcp", "<u4"), ("bwt", "u1")]))
    rec["text"], rec["lcp"], rec["bwt"] = da, lcp, ebwt
    rec["suff"] = rng.integers(0, 1 << 31, size=ebwt.shape[0])
    assert rec.dtype.itemsize == 13
    return rec.tobytes()


@pytest.mark.parametrize("n,tail", [(0, 0), (1, 0), (70_001, 0), (70_001, 7), (200_000, 12)])
def test_gesa_converter_matches_reference(tmp_path, reference, n, tail):
    _need(reference, "EGSAtoBCR")
    rng = np.random.default_rng(n + tail)
    ebwt = rng.integers(0, 256, size=n, dtype=np.uint8)
    lcp = rng.integers(0, 1 << 32, size=n, dtype=np.uint64).astype(np.uint32)
    da = rng.integers(0, 1 << 32, size=n, dtype=np.uint64).astype(np.uint32)
    raw = _gesa_bytes(ebwt, lcp, da, rng) + bytes(rng.integers(0, 256, size=tail, dtype=np.uint8))   # truncated last record
    outs = {}
    for who, exe in (("ref", os.path.join(reference.dir, "EGSAtoBCR")), ("ours", os.path.join(BIN, "EGSAtoBCR"))):
        d = tmp_path / who
        d.mkdir()
        fasta = str(d / "all.fasta")
        with open(fasta + ".4.gesa", "wb") as f:
            f.write(raw)
        r = subprocess.run([exe, fasta, "4"], cwd=str(d), stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True)
        assert r.returncode == 0, r.stderr
        assert f"The total number of elements is {n}" in r.stderr
        outs[who] = [open(fasta + ext, "rb").read() for ext in (".ebwt", ".lcp", ".da")]
    assert outs["ours"] == outs["ref"]
    assert outs["ours"] == [ebwt.tobytes(), lcp.tobytes(), da.tobytes()]
    e, l, d_ = api.read_gesa(str(tmp_path / "ours" / "all.fasta.4.gesa"))
    assert np.array_equal(e, ebwt) and np.array_equal(l, lcp) and np.array_equal(d_, da)
    r = subprocess.run([os.path.join(BIN, "EGSAtoBCR")], stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True)
    assert r.returncode == 1 and "Error usage" in r.stderr and "fastaFile numSeq" in r.stderr


@pytest.mark.parametrize("alpha", [8, 16])
def test_res_txt_writer_matches_reference_bin0(tmp_path, oracle, reference, corpus, alpha):
    _need(reference, "ClusterBWT_DA_bin0")
    done = 0
    for case in corpus:
        n = case["lcp"].shape[0]
        if n < 64:
            continue
        d = tmp_path / f"{case['name']}_{alpha}"
        d.mkdir()
        fasta = str(d / "s+Refs.fasta")
        datagen.write_strand(fasta, case["ebwt"], case["lcp"], case["da"])
        cl, aux = reference.cluster_lcp(fasta, case["n_reads"], case["n_refs"], alpha, 2)
        if aux["maxLen"] > 65536:
            continue
        beta = 0.25 if case["name"] != "wrap" else 0.0
        r = reference._run("ClusterBWT_DA_bin0", [fasta, case["read_len"], beta, 2], str(d))
        assert r.returncode == 0, r.stderr
        sim = oracle.similarity(cl, case["da"], case["ebwt"], case["n_reads"], case["n_refs"])
        b, p = oracle.choose(sim, (case["read_len"] + 1 - alpha) & 0xFFFFFFFF, np.float32(beta))
        ours = str(d / "ours.res.txt")
        api.write_res_txt(ours, b, p)
        assert open(ours, "rb").read() == open(fasta + ".res.txt", "rb").read(), case["name"]
        # reading the text back gives the images with the 5-decimal values
        b2, p2 = api.read_res_txt(ours, case["n_reads"])
        assert np.array_equal(p2, p) and np.array_equal(b2["idRef"], b["idRef"])
        want = np.array([np.float32(float("%.5f" % x)) for x in b["sim"]], dtype=np.float32)
        hdr = np.zeros(b.shape[0], bool)
        hdr[p[p > 0].astype(np.int64)] = True
        assert np.array_equal(b2["sim"], want)
        done += 1
    assert done >= 3


@pytest.mark.parametrize("n_files", [2, 4])
def test_res_txt_reader_then_classify_matches_reference_bin0(tmp_path, oracle, reference, n_files):
    _need(reference, "Classify_bin0")
    rng = np.random.default_rng(90 + n_files)
    n_reads, n_refs, tax_rank = 1500, 12, 1
    lin = str(tmp_path / "lineage.csv")
    with open(lin, "w", newline="") as f:
        f.write(C.lineage_text(C.default_lineage(n_refs, holes=False, seed=3)))
    stems, bins, poss = [], [], []
    for i in range(n_files):
        b, p = _random_rows(rng, n_reads, n_refs)
        stem = str(tmp_path / f"strand{i}.res")
        api.write_res_txt(stem + ".txt", b, p)
        b5, p5 = api.read_res_txt(stem + ".txt", n_reads)
        stems.append(stem); bins.append(b5); poss.append(p5)
    out_ref = str(tmp_path / "ref.csv")
    r = reference._run("Classify_bin0", [n_files] + stems + [n_reads, n_refs, out_ref, lin, tax_rank, 1], str(tmp_path))
    assert r.returncode == 0, r.stderr
    v, m, _ = oracle.fix_rank(lin, tax_rank, n_refs, False)
    typ, tax, sim = oracle.classify(bins, poss, n_refs, v, None, tax_rank, False)
    out_o = str(tmp_path / "oracle.csv")
    oracle.write_csv(out_o, typ, tax, sim)
    assert open(out_ref, "rb").read() == open(out_o, "rb").read()
    assert set(chr(t) for t in typ) >= {"C", "U", "A"}
