#!/usr/bin/env python
"""Generates tests/golden/*.npz by running the UNMODIFIED reference executables (oracle/_ref, built
from /root/reference by oracle/Makefile) on small inputs.  Run in the build container only:

    python tests/golden/make_golden.py

Each .npz holds the inputs (ebwt, lcp, da, parameters) and the reference's output files as raw
bytes: clrs (sorted by pStart), out, res_bin / res_pos for EBWT=1 and EBWT=0; the classify fixtures
hold .res images, the lineage text and the CSV for HIGHER=0/1.  The committed files let the oracle
(CPU tests) and the CUDA path (GPU tests) be checked against the reference itself where
/root/reference does not exist.
"""
import os
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import corpus as C  # noqa: E402
from lime_b200 import datagen  # noqa: E402
from oracle.oracle import Reference  # noqa: E402
from test_oracle_vs_reference import _random_rows  # noqa: E402


def steps12(ref, case, alpha, beta):
    with tempfile.TemporaryDirectory() as d:
        fasta = os.path.join(d, "s+Refs.fasta")
        datagen.write_strand(fasta, case["ebwt"], case["lcp"], case["da"])
        cl, aux = ref.cluster_lcp(fasta, case["n_reads"], case["n_refs"], alpha, 1)
        out = dict(clrs=np.frombuffer(cl.tobytes(), np.uint8), out=np.frombuffer(aux["raw"], np.uint8))
        for ebwt in (1, 0):
            b, p, r = ref.cluster_bwt_da(fasta, case["read_len"], beta, 2, ebwt=bool(ebwt))
            assert b is not None, r.stderr
            out[f"res_bin_{ebwt}"] = np.frombuffer(b.tobytes(), np.uint8)
            out[f"res_pos_{ebwt}"] = np.frombuffer(p.tobytes(), np.uint8)
    return out


def main():
    ref = Reference()
    assert ref.available(), "build oracle/_ref first (make -C oracle ref)"
    cases = [C.exact_case("exact_plain", 11), C.exact_case("exact_iupac_nul", 12, iupac_reads=0.05, iupac_genomes=0.08, terminator=b"\0"),
             C.exact_case("exact_near_dup", 14, genome_lens=(600, 600, 600, 500), near_dup=True),
             C.synth_case("synth_tiny", 40_000, 300, 7, seed=21, p_long=5e-4, max_seg=2000),
             C.wrap_case(n=20000)] + C.edge_cases()[:1]
    for case in cases:
        for alpha in (8, 16):
            beta = 0.0 if case["name"] == "wrap" else 0.25
            g = steps12(ref, case, alpha, beta)
            np.savez_compressed(os.path.join(HERE, f"steps12_{case['name']}_a{alpha}.npz"), ebwt=case["ebwt"], lcp=case["lcp"],
                                da=case["da"], n_reads=case["n_reads"], n_refs=case["n_refs"], read_len=case["read_len"],
                                alpha=alpha, beta=np.float32(beta), **g)
    for n_files in (2, 4):
        rng = np.random.default_rng(1000 + n_files)
        n_reads, n_refs = 600, 12
        lin_txt = C.lineage_text(C.default_lineage(n_refs, holes=True, seed=3), crlf=True, trailing_blank=True)
        rows = [_random_rows(rng, n_reads, n_refs) for _ in range(n_files)]
        fx = dict(n_reads=n_reads, n_refs=n_refs, lineage=np.frombuffer(lin_txt.encode(), np.uint8))
        for i, (b, p) in enumerate(rows):
            fx[f"bin{i}"] = np.frombuffer(b.tobytes(), np.uint8)
            fx[f"pos{i}"] = p
        with tempfile.TemporaryDirectory() as d:
            lin = os.path.join(d, "lineage.csv")
            open(lin, "w", newline="").write(lin_txt)
            stems = []
            for i, (b, p) in enumerate(rows):
                stem = os.path.join(d, f"s{i}.res")
                b.tofile(stem + ".bin"); p.tofile(stem + ".pos"); stems.append(stem)
            for higher in (0, 1):
                for tax_rank in (1, 2):
                    o = os.path.join(d, "o.csv")
                    r = ref.classify(stems, n_reads, n_refs, o, lin, tax_rank, threads=2, higher=bool(higher))
                    assert r.returncode == 0, r.stderr
                    fx[f"csv_h{higher}_t{tax_rank}"] = np.frombuffer(open(o, "rb").read(), np.uint8)
        np.savez_compressed(os.path.join(HERE, f"classify_{n_files}.npz"), **fx)
    print("golden fixtures written to", HERE)


if __name__ == "__main__":
    main()
