#!/usr/bin/env python
"""BASELINE config 1 stand-in: the upstream example cannot be run verbatim -- example/refs.fasta
is missing from the checkout and the preprocessing tools (BCR, egsa, eGap, seqtk) need the network -- so the real
example reads (reads_1.fasta / reads_2.fasta, 10,000 pairs x 100 bp, header = source accession) are classified against
three pseudo-genomes, one per row of example/LineageFile.csv: the concatenation of that accession's mate-1 reads and
reverse-complemented mate-2 reads.  The four strand collections of LiME_paired.sh (mate1 F / RC, mate2 F / RC, each
reads + genomes) are suffix-sorted exactly (datagen.exact_collection_fast, reverse complements made here instead of
`seqtk seq -r`, Preprocessing.sh:55,65) and the UNMODIFIED reference binaries (oracle/_ref) run with the defaults of
LiME_paired.sh:10-12 (alpha=16, beta=0.25, TaxRank=1, threads=1 as in the README quick test).

    python tests/golden/make_config1.py              # every 33rd pair -> tests/golden/config1_example.npz (committed)
    python tests/golden/make_config1.py --pairs all  # all 10,000 pairs: prints the summary for BASELINE.md, writes nothing

Reads of the fourth accession (CP000360, not in the lineage file) are the negative controls.  Run in the build
container only (needs /root/reference)."""
import argparse
import os
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

from lime_b200 import datagen  # noqa: E402
from oracle.oracle import Reference  # noqa: E402

EX = "/root/reference/example"


def read_fasta(path):
    names, seqs = [], []
    for line in open(path):
        line = line.strip()
        if line.startswith(">"):
            names.append(line[1:])
        elif line:
            seqs.append(line)
    return names, seqs


def build(step):
    n1, r1 = read_fasta(os.path.join(EX, "reads_1.fasta"))
    n2, r2 = read_fasta(os.path.join(EX, "reads_2.fasta"))
    assert len(r1) == len(r2) == 10000 and all(a.split("/")[0] == b.split("/")[0] for a, b in zip(n1, n2))
    lineage = open(os.path.join(EX, "LineageFile.csv"), newline="").read()
    accs = [l.split(";")[0] for l in lineage.splitlines()[1:] if l.strip()]
    sel = list(range(0, len(r1), step))
    r1, r2, acc_of = [r1[i] for i in sel], [r2[i] for i in sel], [n1[i].split("-")[0] for i in sel]
    genomes = ["".join(a for a, x in zip(r1, acc_of) if x == acc) + "".join(datagen.revcomp(b) for b, x in zip(r2, acc_of) if x == acc)
               for acc in accs]
    strands = [r1, [datagen.revcomp(r) for r in r1], r2, [datagen.revcomp(r) for r in r2]]     # LiME_paired.sh:15-18
    return strands, genomes, lineage, accs, acc_of


def run_reference(strands, genomes, lineage, alpha=16, beta=0.25, tax_rank=1, read_len=100, threads=1):
    ref = Reference()
    assert ref.available(), "build oracle/_ref first (make -C oracle ref)"
    n_reads, n_refs = len(strands[0]), len(genomes)
    out = {}
    with tempfile.TemporaryDirectory() as d:
        lin = os.path.join(d, "LineageFile.csv")
        open(lin, "w", newline="").write(lineage)
        fastas = []
        for s, reads in enumerate(strands):
            e, l, da = datagen.exact_collection_fast(reads, genomes)
            fasta = os.path.join(d, f"strand{s}+Refs.fasta")
            datagen.write_strand(fasta, e, l, da)
            fastas.append(fasta)
            cl, aux = ref.cluster_lcp(fasta, n_reads, n_refs, alpha, threads)
            b, p, r = ref.cluster_bwt_da(fasta, read_len, beta, threads, ebwt=True)
            assert b is not None, r.stderr
            out[s] = dict(ebwt=e, lcp=l, da=da, out=aux["raw"], clrs=cl.tobytes(), res_bin=b.tobytes(), res_pos=p.tobytes())
        csv = os.path.join(d, "RESULTS.csv")
        r = ref.classify([f + ".res" for f in fastas], n_reads, n_refs, csv, lin, tax_rank, threads)
        assert r.returncode == 0, r.stderr
        out["csv"] = open(csv, "rb").read()
    return out


def summary(out, lineage, accs, acc_of):
    tax = {l.split(";")[0]: l.split(";")[1] for l in lineage.splitlines()[1:] if l.strip()}
    rows = [l.split(",") for l in out["csv"].decode().splitlines()[1:]]
    kinds, right, wrong, neg_u, neg = {}, 0, 0, 0, 0
    for row, acc in zip(rows, acc_of):
        kinds[row[0]] = kinds.get(row[0], 0) + 1
        if acc in tax:
            right += row[0] == "C" and row[2] == tax[acc]
            wrong += row[0] == "C" and row[2] != tax[acc]
        else:
            neg += 1
            neg_u += row[0] == "U"
    return dict(reads=len(rows), kinds=kinds, correct_species=right, wrong_species=wrong, negatives=neg, negatives_unclassified=neg_u,
                positions_per_strand=[int(out[s]["lcp"].shape[0]) for s in range(4)])


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--pairs", default="33", help="take every k-th pair (default 33), or 'all'")
    a = ap.parse_args()
    step = 1 if a.pairs == "all" else int(a.pairs)
    strands, genomes, lineage, accs, acc_of = build(step)
    out = run_reference(strands, genomes, lineage)
    print(summary(out, lineage, accs, acc_of))
    if step == 1:
        return
    arrs = {"csv": np.frombuffer(out["csv"], np.uint8), "lineage": np.frombuffer(lineage.encode(), np.uint8),
            "n_reads": len(strands[0]), "n_refs": len(genomes), "acc_of": np.array(acc_of)}
    for s in range(4):
        for k, v in out[s].items():
            arrs[f"{k}_{s}"] = v if isinstance(v, np.ndarray) else np.frombuffer(v, np.uint8)
    np.savez_compressed(os.path.join(HERE, "config1_example.npz"), **arrs)
    print("wrote", os.path.join(HERE, "config1_example.npz"), os.path.getsize(os.path.join(HERE, "config1_example.npz")), "bytes")


if __name__ == "__main__":
    main()
