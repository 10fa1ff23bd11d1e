"""Shared test corpus: small inputs that exercise every quirk of SURVEY.md Appendix A.

A case is a dict: name, ebwt (u8), lcp (u32), da (u32), n_reads, n_refs, read_len, lineage (text),
optionally mates (for paired cases the four strands are separate cases sharing n_reads).
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

from lime_b200 import datagen  # noqa: E402

HEADER = "Acc;Species_TaxID;Genus_TaxID;Family_TaxID;Order_TaxID;Class_TaxID;Phylum_TaxID"


def lineage_text(rows, crlf=False, trailing_blank=False, final_newline=True):
    eol = "\r\n" if crlf else "\n"
    lines = [HEADER] + [";".join(str(x) for x in r) for r in rows]
    txt = eol.join(lines)
    if final_newline:
        txt += eol
    if trailing_blank:
        txt += eol
    return txt


def default_lineage(n_refs, groups=3, seed=0, holes=False):
    """n_refs rows; species ids repeat so that several genomes share a taxon; upper ranks nest."""
    rng = np.random.default_rng(seed)
    rows = []
    for g in range(n_refs):
        sp = 100 + (g % max(1, (n_refs + 1) // 2))
        ge = 1000 + sp % groups
        fa = 2000 + ge % 2
        row = [f"ACC{g}", sp, ge, fa, 3000, 4000, 5000]
        if holes and rng.random() < 0.2:
            row[4] = ""                      # empty order field (stays 0 in m_corRef)
        if holes and rng.random() < 0.3:
            row[3] = 0                       # zero family: never assigned at that rank (A20)
        rows.append(row)
    return rows


def exact_case(name, seed, n_reads=60, read_len=50, genome_lens=(600, 500, 700), err=0.02, iupac_reads=0.0,
               iupac_genomes=0.0, terminator=b"$", near_dup=False, strand="F"):
    rng = np.random.default_rng(seed)
    genomes = []
    for i, L in enumerate(genome_lens):
        if near_dup and i > 0:
            g = list(genomes[0][:L].ljust(L, "A"))
            for j in range(L):
                if rng.random() < 0.01:
                    g[j] = "ACGT"[int(rng.integers(4))]
            genomes.append("".join(g))
        else:
            genomes.append(datagen.random_dna(rng, L))
    if iupac_genomes > 0:
        gg = []
        for g in genomes:
            g = list(g)
            for j in range(len(g)):
                if rng.random() < iupac_genomes:
                    g[j] = "RYSWKMBDHVN"[int(rng.integers(11))]
            gg.append("".join(g))
        genomes = gg
    reads, src = datagen.sample_reads(rng, genomes, n_reads, read_len, err=err, iupac=iupac_reads)
    if strand == "RC":
        reads = [datagen.revcomp(r) for r in reads]
    ebwt, lcp, da = datagen.exact_collection(reads, genomes, terminator)
    return dict(name=name, ebwt=ebwt, lcp=lcp, da=da, n_reads=n_reads, n_refs=len(genome_lens),
                read_len=read_len, reads=reads, genomes=genomes)


def synth_case(name, n, n_reads, n_refs, read_len=100, seed=1, **kw):
    ebwt, lcp, da = datagen.synth_arrays(n, n_reads, n_refs, read_len, seed=seed, **kw)
    return dict(name=name, ebwt=ebwt.numpy(), lcp=lcp.numpy().view(np.uint32), da=da.numpy().view(np.uint32),
                n_reads=n_reads, n_refs=n_refs, read_len=read_len)


def wrap_case(name="wrap", n=60000, seed=5):
    """Few reads and genomes, long clusters: cell sums far above 255 -> uint8 wrap (A10, P11),
    read counts >= 256 inside one cluster (A7 wrap) and genome counts saturating at 255 (A9)."""
    rng = np.random.default_rng(seed)
    n_reads, n_refs = 5, 2
    lcp = np.where(rng.random(n) < 0.002, 0, 20).astype(np.uint32)
    lcp[0] = 0
    da = rng.integers(0, n_reads + n_refs, size=n).astype(np.uint32)
    ebwt = np.frombuffer(bytes(rng.choice(list(b"ACGTN$RY"), size=n).tolist()), dtype=np.uint8).copy()
    return dict(name=name, ebwt=ebwt, lcp=lcp, da=da, n_reads=n_reads, n_refs=n_refs, read_len=100)


def edge_cases():
    """Hand-made arrays for the boundary rules of step 1 (A1-A3)."""
    cases = []
    # lcp[0] >= alpha: leading run is skipped (A2); cluster open at end of input (A1)
    lcp = np.array([20, 20, 20, 3, 20, 20, 1, 0, 16, 15, 16, 16, 2, 20], dtype=np.uint32)
    da = np.array([0, 5, 1, 0, 5, 6, 1, 2, 6, 0, 1, 5, 1, 6], dtype=np.uint32)
    ebwt = np.frombuffer(b"ACGTACGTACGT$N", dtype=np.uint8).copy()
    cases.append(dict(name="edge_leading_open_end", ebwt=ebwt, lcp=lcp, da=da, n_reads=5, n_refs=3, read_len=30))
    # no position below alpha at all -> no clusters
    cases.append(dict(name="edge_all_high", ebwt=np.full(50, 65, np.uint8), lcp=np.full(50, 99, np.uint32),
                      da=(np.arange(50) % 7).astype(np.uint32), n_reads=4, n_refs=3, read_len=30))
    # all singletons
    cases.append(dict(name="edge_all_low", ebwt=np.full(50, 67, np.uint8), lcp=np.zeros(50, np.uint32),
                      da=(np.arange(50) % 7).astype(np.uint32), n_reads=4, n_refs=3, read_len=30))
    # one position / empty
    cases.append(dict(name="edge_one", ebwt=np.array([65], np.uint8), lcp=np.array([0], np.uint32),
                      da=np.array([0], np.uint32), n_reads=1, n_refs=1, read_len=30))
    cases.append(dict(name="edge_empty", ebwt=np.zeros(0, np.uint8), lcp=np.zeros(0, np.uint32),
                      da=np.zeros(0, np.uint32), n_reads=1, n_refs=1, read_len=30))
    return cases


def small_corpus():
    """Cases cheap enough for every test run (oracle and reference finish in seconds)."""
    c = []
    c.append(exact_case("exact_plain", 11))
    c.append(exact_case("exact_iupac_nul", 12, iupac_reads=0.05, iupac_genomes=0.08, terminator=b"\0"))
    c.append(exact_case("exact_hash_term", 13, iupac_reads=0.02, iupac_genomes=0.02, terminator=b"#"))
    c.append(exact_case("exact_near_dup", 14, genome_lens=(600, 600, 600, 500), near_dup=True))
    c.append(synth_case("synth_small", 200_000, 1000, 7, seed=21, p_long=2e-4, max_seg=5000))
    c.append(synth_case("synth_many_refs", 300_000, 800, 133, seed=22, p_long=1e-4, max_seg=3000))
    c.append(synth_case("synth_trlcp", 150_000, 500, 5, seed=23, trlcp=32, p_long=0))
    c.append(wrap_case())
    c.extend(edge_cases())
    return c
