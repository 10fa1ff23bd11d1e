"""Pins the oracle (oracle/lime_oracle.c) against the UNMODIFIED reference executables built into
oracle/_ref/ by oracle/Makefile: same inputs, files diffed (SURVEY.md section 8c comparison rules)."""
import os

import numpy as np
import pytest

import corpus as C
from lime_b200 import datagen


def _write_case(tmp, case, fname="s+Refs.fasta"):
    fasta = os.path.join(tmp, fname)
    datagen.write_strand(fasta, case["ebwt"], case["lcp"], case["da"])
    return fasta


@pytest.mark.parametrize("alpha", [8, 16])
@pytest.mark.parametrize("threads", [1, 3])
def test_step1_and_2_match_reference(tmp_path, oracle, reference, corpus, alpha, threads):
    for case in corpus:
        n = case["lcp"].shape[0]
        if n < 8 * threads or n == 0:       # the reference itself misbehaves on empty / tiny chunks
            continue
        d = tmp_path / f"{case['name']}_{alpha}_{threads}"
        d.mkdir()
        fasta = _write_case(str(d), case)
        cl_ref, aux = reference.cluster_lcp(fasta, case["n_reads"], case["n_refs"], alpha, threads)
        cl, max_len = oracle.cluster_detect(case["lcp"], case["da"], case["n_reads"], alpha)
        assert aux["nClusters"] == cl.shape[0], case["name"]
        assert aux["maxLen"] == max_len, case["name"]
        assert np.array_equal(cl_ref, cl), case["name"]
        if max_len > 65536:
            continue
        for ebwt in (True, False):
            beta = 0.25 if case["name"] != "wrap" else 0.0
            bin_ref, pos_ref, r = reference.cluster_bwt_da(fasta, case["read_len"], beta, threads, ebwt=ebwt)
            assert bin_ref is not None, r.stderr
            sim = oracle.similarity(cl, case["da"], case["ebwt"] if ebwt else None, case["n_reads"], case["n_refs"])
            norm = (case["read_len"] + 1 - alpha) & 0xFFFFFFFF
            bin_o, pos_o = oracle.choose(sim, norm, np.float32(beta))
            assert bin_ref.tobytes() == bin_o.tobytes(), (case["name"], ebwt)
            assert pos_ref.tobytes() == pos_o.tobytes(), (case["name"], ebwt)


def _random_rows(rng, n_reads, n_refs, p_empty=0.4, near_tie=True):
    """Random .res.bin/.res.pos images with many near-ties (SURVEY probe P13)."""
    recs = [(0.0, 0)]
    pos = np.zeros(n_reads, dtype=np.uint64)
    for r in range(n_reads):
        if rng.random() < p_empty:
            continue
        k = int(rng.integers(1, min(n_refs, 6) + 1))
        ids = np.sort(rng.choice(n_refs, size=k, replace=False))
        base = rng.integers(20, 86) / 85.0
        sims = []
        for _ in range(k):
            if near_tie and rng.random() < 0.6:
                sims.append(np.float32(max(1, int(base * 85) - int(rng.integers(0, 3))) / np.float32(85)))
            else:
                sims.append(np.float32(rng.integers(1, 86)) / np.float32(85))
        pos[r] = len(recs)
        recs.append((max(sims), k))
        recs.extend(zip(sims, ids))
    arr = np.zeros(len(recs), dtype=[("sim", "<f4"), ("idRef", "<u4")])
    arr["sim"] = [x[0] for x in recs]
    arr["idRef"] = [x[1] for x in recs]
    return arr, pos


@pytest.mark.parametrize("n_files", [2, 4])
@pytest.mark.parametrize("higher", [False, True])
@pytest.mark.parametrize("tax_rank", [1, 2])
def test_step3_matches_reference(tmp_path, oracle, reference, n_files, higher, tax_rank):
    rng = np.random.default_rng(100 + n_files + 10 * higher + tax_rank)
    n_reads, n_refs = 1500, 12
    lin = os.path.join(tmp_path, "lineage.csv")
    with open(lin, "w", newline="") as f:
        f.write(C.lineage_text(C.default_lineage(n_refs, holes=True, seed=3), crlf=True, trailing_blank=True))
    stems, bins, poss = [], [], []
    for i in range(n_files):
        b, p = _random_rows(rng, n_reads, n_refs)
        stem = os.path.join(tmp_path, f"strand{i}.res")
        b.tofile(stem + ".bin")
        p.tofile(stem + ".pos")
        stems.append(stem); bins.append(b); poss.append(p)
    out_ref = os.path.join(tmp_path, "ref.csv")
    r = reference.classify(stems, n_reads, n_refs, out_ref, lin, tax_rank, threads=2, higher=higher)
    assert r.returncode == 0, r.stderr
    v, m, found = oracle.fix_rank(lin, tax_rank, n_refs, higher)
    assert v is not None
    typ, tax, sim = oracle.classify(bins, poss, n_refs, v, m, tax_rank, higher)
    out_o = os.path.join(tmp_path, "oracle.csv")
    oracle.write_csv(out_o, typ, tax, sim)
    assert open(out_ref, "rb").read() == open(out_o, "rb").read()
    kinds = set(chr(t) for t in typ)
    assert "C" in kinds and "U" in kinds and (("H" in kinds) if higher else ("A" in kinds))


@pytest.mark.parametrize("shape", ["cfg3_930genomes_ebwt0", "cfg4_1000genomes_higher1_trlcp32", "cfg5_alpha12", "cfg5_alpha20", "cfg5_alpha24"])
def test_baseline_config_shapes_match_reference(tmp_path, oracle, reference, shape):
    """Whole jobs in the shapes of BASELINE configs 3, 4 and 5 (reduced sizes), the three reference executables chained
    through their files against the oracle: .clrs, .res.bin / .res.pos of every strand and the CSV, byte for byte."""
    from lime_b200.pipeline import synthetic_lineage
    if shape.startswith("cfg3"):
        n, n_reads, n_refs, read_len, alpha, ebwt, higher, n_str, kw = 260_000, 1500, 930, 100, 16, False, False, 4, {}
    elif shape.startswith("cfg4"):
        n, n_reads, n_refs, read_len, alpha, ebwt, higher, n_str, kw = 260_000, 1500, 1000, 150, 16, True, True, 4, {"trlcp": 32}
    else:
        n, n_reads, n_refs, read_len, alpha, ebwt, higher, n_str, kw = 200_000, 3000, 50, 100, int(shape[10:]), True, False, 2, {}
    rows, v, m = synthetic_lineage(n_refs)
    lin = os.path.join(tmp_path, "lineage.csv")
    with open(lin, "w", newline="") as f:
        f.write(C.lineage_text(rows))
    vo, mo, found = oracle.fix_rank(lin, 1, n_refs, higher)
    assert found == n_refs and np.array_equal(vo, v)
    stems, bins, poss = [], [], []
    for s_ in range(n_str):
        case = C.synth_case(f"{shape}_{s_}", n + 17 * s_, n_reads, n_refs, read_len=read_len, seed=800 + s_, p_long=2e-5, max_seg=20000, **kw)
        d = tmp_path / f"strand{s_}"
        d.mkdir()
        fasta = _write_case(str(d), case)
        cl_ref, aux = reference.cluster_lcp(fasta, n_reads, n_refs, alpha, 2)
        cl, max_len = oracle.cluster_detect(case["lcp"], case["da"], n_reads, alpha)
        assert aux["nClusters"] == cl.shape[0] and aux["maxLen"] == max_len and np.array_equal(cl_ref, cl)
        assert max_len <= 65536
        bin_ref, pos_ref, r = reference.cluster_bwt_da(fasta, read_len, 0.25, 2, ebwt=ebwt)
        assert bin_ref is not None, r.stderr
        sim = oracle.similarity(cl, case["da"], case["ebwt"] if ebwt else None, n_reads, n_refs)
        b, p_ = oracle.choose(sim, (read_len + 1 - alpha) & 0xFFFFFFFF, np.float32(0.25))
        assert bin_ref.tobytes() == b.tobytes() and pos_ref.tobytes() == p_.tobytes()
        stem = os.path.join(tmp_path, f"s{s_}.res")
        b.tofile(stem + ".bin"); p_.tofile(stem + ".pos")
        stems.append(stem); bins.append(b); poss.append(p_)
    out_ref = os.path.join(tmp_path, "ref.csv")
    r = reference.classify(stems, n_reads, n_refs, out_ref, lin, 1, threads=2, higher=higher)
    assert r.returncode == 0, r.stderr
    typ, tax, simc = oracle.classify(bins, poss, n_refs, vo, mo, 1, higher)
    out_o = os.path.join(tmp_path, "oracle.csv")
    oracle.write_csv(out_o, typ, tax, simc)
    assert open(out_ref, "rb").read() == open(out_o, "rb").read()
    assert "C" in set(chr(t) for t in typ)


def test_lineage_without_final_newline_is_rejected_like_reference(tmp_path, oracle, reference):
    lin = os.path.join(tmp_path, "lineage.csv")
    with open(lin, "w", newline="") as f:
        f.write(C.lineage_text(C.default_lineage(3), final_newline=False))
    v, m, found = oracle.fix_rank(lin, 1, 3)
    assert v is None and found == 2
    b, p = _random_rows(np.random.default_rng(1), 10, 3)
    stems = []
    for i in range(2):
        stem = os.path.join(tmp_path, f"s{i}.res")
        b.tofile(stem + ".bin"); p.tofile(stem + ".pos"); stems.append(stem)
    r = reference.classify(stems, 10, 3, os.path.join(tmp_path, "o.csv"), lin, 1)
    assert r.returncode == 1 and "Number of taxIDs = 2" in r.stderr
