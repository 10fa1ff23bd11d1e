"""The drop-in executables (lime_b200/host/bin) against the oracle and, where built, against the
unmodified reference executables: same argv, same files, byte-identical outputs."""
import os
import subprocess

import numpy as np
import pytest

import corpus as C
from lime_b200 import datagen
from test_oracle_vs_reference import _random_rows

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BIN = os.path.join(ROOT, "lime_b200", "host", "bin")


def _run(exe, args, cwd):
    return subprocess.run([os.path.join(BIN, exe)] + [str(a) for a in args], cwd=cwd,
                          stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True)


def test_executables_are_built_and_print_the_reference_usage():
    for exe, n in (("ClusterLCP", "fileFasta numReads numGenomes alpha threads"),
                   ("ClusterBWT_DA", "fileFasta readLen beta threads"),
                   ("ClusterBWT_DA_ebwt0", "fileFasta readLen beta threads"),
                   ("Classify", "N fileInput1 fileInput2 ... fileInputN numReads numGenomes fileOutput fileTaxo taxRank numThreads"),
                   ("Classify_higher1", "N fileInput1"),
                   ("LiME_fused", "N File_1 ... File_N output numReads numGenomes LineageFile readLen threads"),
                   ("LiME_fused_ebwt0", "N File_1"), ("LiME_fused_higher1", "N File_1"),
                   ("ClusterBWT_DA_bin0", "fileFasta readLen beta threads"), ("Classify_bin0", "N fileInput1")):
        r = _run(exe, [], ROOT)
        assert r.returncode == 1 and "Error usage" in r.stderr and n in r.stderr, (exe, r.stderr)
    r = _run("Classify", [3, "a", "b", "c", 1, 1, "o", "t", 1, 1], ROOT)
    assert r.returncode == 1 and "allowed number of input files is 2" in r.stderr
    r = _run("LiME_fused", [3, "a", "b", "c", "o", 1, 1, "t", 100, 1], ROOT)
    assert r.returncode == 1 and "allowed number of input files is 2" in r.stderr


@pytest.mark.gpu
@pytest.mark.parametrize("alpha", [8, 16])
def test_steps_1_2_cli_files_match(tmp_path, oracle, corpus, alpha):
    for case in corpus:
        n = case["lcp"].shape[0]
        if n == 0:
            continue
        d = tmp_path / f"{case['name']}_{alpha}"
        d.mkdir()
        fasta = str(d / "s+Refs.fasta")
        datagen.write_strand(fasta, case["ebwt"], case["lcp"], case["da"])
        r = _run("ClusterLCP", [fasta, case["n_reads"], case["n_refs"], alpha, 4], str(d))
        assert r.returncode == 0, r.stderr
        cl_o, max_o = oracle.cluster_detect(case["lcp"], case["da"], case["n_reads"], alpha)
        assert open(f"{fasta}.{alpha}.clrs", "rb").read() == cl_o.tobytes(), case["name"]
        aux = open(str(d / "s+Refs.out"), "rb").read()
        exp = (np.array([case["n_reads"], case["n_refs"], alpha], "<u4").tobytes() +
               np.array([max_o, cl_o.shape[0]], "<u8").tobytes())
        assert aux == exp, case["name"]
        assert f"Total number of clusters: {cl_o.shape[0]}." in r.stdout
        for exe, ebwt in (("ClusterBWT_DA", True), ("ClusterBWT_DA_ebwt0", False)):
            beta = 0.0 if case["name"] == "wrap" else 0.25
            r = _run(exe, [fasta, case["read_len"], beta, 4], str(d))
            if max_o > 65536:
                assert r.returncode == 1 and "greater than sizeMaxBuf" in r.stderr
                continue
            assert r.returncode == 0, r.stderr
            sim = oracle.similarity(cl_o, case["da"], case["ebwt"] if ebwt else None, case["n_reads"], case["n_refs"])
            b, p = oracle.choose(sim, (case["read_len"] + 1 - alpha) & 0xFFFFFFFF, np.float32(beta))
            assert open(fasta + ".res.bin", "rb").read() == b.tobytes(), (case["name"], exe)
            assert open(fasta + ".res.pos", "rb").read() == p.tobytes(), (case["name"], exe)


@pytest.mark.gpu
@pytest.mark.parametrize("n_files", [2, 4])
@pytest.mark.parametrize("higher", [False, True])
def test_classify_cli_csv_matches(tmp_path, oracle, n_files, higher):
    rng = np.random.default_rng(50 + n_files + 10 * higher)
    n_reads, n_refs, tax_rank = 2000, 12, 1
    lin = str(tmp_path / "lineage.csv")
    with open(lin, "w", newline="") as f:
        f.write(C.lineage_text(C.default_lineage(n_refs, holes=True, seed=3), crlf=True, trailing_blank=True))
    stems, bins, poss = [], [], []
    for i in range(n_files):
        b, p = _random_rows(rng, n_reads, n_refs)
        stem = str(tmp_path / f"strand{i}.res")
        b.tofile(stem + ".bin"); p.tofile(stem + ".pos")
        stems.append(stem); bins.append(b); poss.append(p)
    out = str(tmp_path / "ours.csv")
    r = _run("Classify_higher1" if higher else "Classify", [n_files] + stems + [n_reads, n_refs, out, lin, tax_rank, 2],
             str(tmp_path))
    assert r.returncode == 0, r.stderr
    v, m, _ = oracle.fix_rank(lin, tax_rank, n_refs, higher)
    typ, tax, sim = oracle.classify(bins, poss, n_refs, v, m, tax_rank, higher)
    exp = str(tmp_path / "oracle.csv")
    oracle.write_csv(exp, typ, tax, sim)
    assert open(out, "rb").read() == open(exp, "rb").read()
    assert f"Number of successfully classified reads: {int((typ == ord('C')).sum())}/{n_reads};" in r.stdout
    # lineage without a final newline loses its last row -> the reference's size-mismatch exit (A15)
    with open(lin, "w", newline="") as f:
        f.write(C.lineage_text(C.default_lineage(n_refs), final_newline=False))
    r = _run("Classify", [n_files] + stems + [n_reads, n_refs, out, lin, tax_rank, 2], str(tmp_path))
    assert r.returncode == 1 and f"Number of taxIDs = {n_refs - 1}" in r.stderr


@pytest.mark.gpu
@pytest.mark.parametrize("n_files,exe,ebwt,higher", [(4, "LiME_fused", True, False), (2, "LiME_fused_ebwt0", False, False),
                                                    (4, "LiME_fused_higher1", True, True)])
def test_fused_driver_equals_the_three_steps(tmp_path, oracle, n_files, exe, ebwt, higher):
    """LiME_fused (LiME_paired.sh in one process) against the chain ClusterLCP -> ClusterBWT_DA -> Classify
    of the drop-in executables and against the oracle: same CSV; with --keep the same intermediate files."""
    alpha, beta, tax_rank, read_len = 16, 0.25, 1, 100
    n_reads, n_refs = 3000, 12
    lin = str(tmp_path / "lineage.csv")
    with open(lin, "w", newline="") as f:
        f.write(C.lineage_text(C.default_lineage(n_refs, holes=False, seed=5)))
    fastas, bins, poss, clusters = [], [], [], []
    for i in range(n_files):
        case = C.synth_case(f"fz{i}", 400_000 + 37 * i, n_reads, n_refs, seed=300 + i, p_long=2e-4, max_seg=3000)
        fasta = str(tmp_path / f"mate{i}+Refs.fasta")
        datagen.write_strand(fasta, case["ebwt"], case["lcp"], case["da"])
        fastas.append(fasta)
        cl, mx = oracle.cluster_detect(case["lcp"], case["da"], n_reads, alpha)
        clusters.append((cl, mx))
        sim = oracle.similarity(cl, case["da"], case["ebwt"] if ebwt else None, n_reads, n_refs)
        b, p = oracle.choose(sim, read_len + 1 - alpha, np.float32(beta))
        bins.append(b); poss.append(p)
    v, m, _ = oracle.fix_rank(lin, tax_rank, n_refs, higher)
    typ, tax, sim = oracle.classify(bins, poss, n_refs, v, m, tax_rank, higher)
    exp = str(tmp_path / "oracle.csv")
    oracle.write_csv(exp, typ, tax, sim)
    # fused, nothing kept
    out = str(tmp_path / "fused.csv")
    r = _run(exe, [n_files] + fastas + [out, n_reads, n_refs, lin, read_len, 8, alpha, beta, tax_rank], str(tmp_path))
    assert r.returncode == 0, r.stderr
    assert open(out, "rb").read() == open(exp, "rb").read()
    assert not os.path.exists(fastas[0] + ".res.bin") and not os.path.exists(f"{fastas[0]}.{alpha}.clrs")
    assert f"Number of successfully classified reads: {int((typ == ord('C')).sum())}/{n_reads};" in r.stdout
    # fused with --keep (default alpha / beta / TaxRank of LiME_paired.sh) vs the three executables
    out2 = str(tmp_path / "fused_keep.csv")
    r = _run(exe, [n_files] + fastas + [out2, n_reads, n_refs, lin, read_len, 8, "--keep"], str(tmp_path))
    assert r.returncode == 0, r.stderr
    assert open(out2, "rb").read() == open(exp, "rb").read()
    # the kept files are what ClusterLCP / ClusterBWT_DA write: checked against the oracle ...
    for i, fa in enumerate(fastas):
        assert open(fa + ".res.bin", "rb").read() == bins[i].tobytes()
        assert open(fa + ".res.pos", "rb").read() == poss[i].tobytes()
        assert open(f"{fa}.{alpha}.clrs", "rb").read() == clusters[i][0].tobytes()
        aux = open(fa[:fa.find(".fasta")] + ".out", "rb").read()
        assert aux == (np.array([n_reads, n_refs, alpha], "<u4").tobytes() +
                       np.array([clusters[i][1], clusters[i][0].shape[0]], "<u8").tobytes())
    if n_files != 2:
        return
    # ... and, for the single-end case, against the chain of the three drop-in executables
    kept = {}
    for fa in fastas:
        for suffix in (f".{alpha}.clrs", ".res.bin", ".res.pos"):
            kept[fa + suffix] = open(fa + suffix, "rb").read()
            os.remove(fa + suffix)
        aux = fa[:fa.find(".fasta")] + ".out"
        kept[aux] = open(aux, "rb").read()
        os.remove(aux)
    for i, fa in enumerate(fastas):
        assert _run("ClusterLCP", [fa, n_reads, n_refs, alpha, 4], str(tmp_path)).returncode == 0
        assert _run("ClusterBWT_DA" if ebwt else "ClusterBWT_DA_ebwt0", [fa, read_len, beta, 4], str(tmp_path)).returncode == 0
    for path, data in kept.items():
        assert open(path, "rb").read() == data, path
    out3 = str(tmp_path / "chain.csv")
    r = _run("Classify_higher1" if higher else "Classify", [n_files] + [fa + ".res" for fa in fastas] + [n_reads, n_refs, out3, lin, tax_rank, 2],
             str(tmp_path))
    assert r.returncode == 0, r.stderr
    assert open(out3, "rb").read() == open(exp, "rb").read()


@pytest.mark.gpu
def test_bin0_text_variants(tmp_path, oracle):
    """BIN=0 builds: ClusterBWT_DA_bin0 writes <fasta>.res.txt, Classify_bin0 reads <in>.txt
    (the text format itself is pinned against the reference in tests/test_formats.py)."""
    from lime_b200 import api
    alpha, beta, read_len, n_reads, n_refs, tax_rank = 16, 0.25, 100, 2500, 12, 1
    lin = str(tmp_path / "lineage.csv")
    with open(lin, "w", newline="") as f:
        f.write(C.lineage_text(C.default_lineage(n_refs, holes=False, seed=5)))
    stems, bins, poss = [], [], []
    for i in range(2):
        case = C.synth_case(f"tx{i}", 300_000 + 11 * i, n_reads, n_refs, seed=400 + i, p_long=2e-4, max_seg=3000)
        fasta = str(tmp_path / f"m{i}+Refs.fasta")
        datagen.write_strand(fasta, case["ebwt"], case["lcp"], case["da"])
        assert _run("ClusterLCP", [fasta, n_reads, n_refs, alpha, 4], str(tmp_path)).returncode == 0
        r = _run("ClusterBWT_DA_bin0", [fasta, read_len, beta, 4], str(tmp_path))
        assert r.returncode == 0, r.stderr
        assert not os.path.exists(fasta + ".res.bin")
        cl, _ = oracle.cluster_detect(case["lcp"], case["da"], n_reads, alpha)
        b, p = oracle.choose(oracle.similarity(cl, case["da"], case["ebwt"], n_reads, n_refs), read_len + 1 - alpha, np.float32(beta))
        exp_txt = str(tmp_path / f"exp{i}.txt")
        api.write_res_txt(exp_txt, b, p)
        assert open(fasta + ".res.txt", "rb").read() == open(exp_txt, "rb").read()
        b5, p5 = api.read_res_txt(exp_txt, n_reads)
        stems.append(fasta + ".res"); bins.append(b5); poss.append(p5)
    out = str(tmp_path / "ours.csv")
    r = _run("Classify_bin0", [2] + stems + [n_reads, n_refs, out, lin, tax_rank, 2], str(tmp_path))
    assert r.returncode == 0, r.stderr
    v, m, _ = oracle.fix_rank(lin, tax_rank, n_refs, False)
    typ, tax, sim = oracle.classify(bins, poss, n_refs, v, None, tax_rank, False)
    exp = str(tmp_path / "oracle.csv")
    oracle.write_csv(exp, typ, tax, sim)
    assert open(out, "rb").read() == open(exp, "rb").read()
    r = _run("Classify_bin0", [2, str(tmp_path / "nope"), stems[1], n_reads, n_refs, out, lin, tax_rank, 2], str(tmp_path))
    assert r.returncode == 1 and "Error opening" in r.stderr
