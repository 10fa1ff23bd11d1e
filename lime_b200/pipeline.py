"""The strand loop of the LiME drivers on one GPU, on top of the C ABI.

LiME_paired.sh:44-79 runs ClusterLCP and ClusterBWT_DA once per strand file (mate1 F, mate1 RC,
mate2 F, mate2 RC -- or F, RC for single-end reads) and then one Classify over the 2 or 4 results.
Here the per-strand results stay in device memory (strand slots of the context) and Classify
reads them there; nothing is written to disk unless the caller copies a slot out.
"""
import numpy as np

from ._lib import MEM_HOST, MEM_DEVICE, ERR_RETRY, LimeError  # noqa: F401


def norm_of(read_len, alpha):
    """norm = readLen + 1 - minLCP (ClusterBWT_DA.cpp:555); readLen is a uchar there (:519-521)."""
    return ((int(read_len) & 0xFF) + 1 - int(alpha)) & 0xFFFFFFFF


def run_strand(ctx, slot, strand, n_reads, n_refs, read_len, alpha, beta, ebwt_mode, mem=MEM_HOST,
               fused=True, keep_clusters=False):
    """Steps 1 and 2 of one strand into strand slot `slot` (fused: one pass over the inputs,
    the cluster list is not materialised unless keep_clusters).

    strand: dict with lcp, da, ebwt (numpy arrays for MEM_HOST, device pointers for MEM_DEVICE)
    and n (positions).  Returns (n_clusters, max_len, n_records)."""
    ebwt = strand.get("ebwt") if ebwt_mode else None
    ctx.load(strand["lcp"], strand["da"], ebwt, n_reads, n_refs, mem=mem, n_total=strand["n"])
    if fused:
        n_cl, max_len = ctx.cluster_similarity(alpha, ebwt_mode, keep_clusters)
    else:
        n_cl, max_len = ctx.cluster_detect(alpha)
        ctx.similarity(ebwt_mode)
    n_rec = ctx.choose(slot, norm_of(read_len, alpha), beta)
    return n_cl, max_len, n_rec


def run_job(ctx, strands, n_reads, n_refs, read_len, alpha, beta, ebwt_mode, v_corRef, m_corRef=None,
            tax_rank=1, higher=False, mem=MEM_HOST, fused=True, deferred=None):
    """A whole single-end (2 strands) or paired-end (4 strands) job: steps 1-2 per strand, then
    step 3.  Returns ((type u8, taxid u32, sim f32) per read, per-strand stats).

    deferred (default: with the fused path): the strands are queued without a host synchronisation
    (lime_strand_run) and checked by lime_classify; LIME_ERR_RETRY redoes the job with the
    synchronising calls (exact sizes, exact uint8-wrap path)."""
    assert len(strands) in (2, 4)            # Classify.cpp:334-338
    if deferred is None:
        deferred = fused
    if deferred and fused:
        try:
            for s, st in enumerate(strands):
                ebwt = st.get("ebwt") if ebwt_mode else None
                ctx.load(st["lcp"], st["da"], ebwt, n_reads, n_refs, mem=mem, n_total=st["n"])
                ctx.strand_run(s, alpha, ebwt_mode, norm_of(read_len, alpha), beta)
            out = ctx.classify(len(strands), n_refs, v_corRef, m_corRef, tax_rank, higher)
            return out, [ctx.strand_status(s)[:3] for s in range(len(strands))]
        except LimeError as e:
            if e.code != ERR_RETRY:
                raise
    stats = [run_strand(ctx, s, st, n_reads, n_refs, read_len, alpha, beta, ebwt_mode, mem, fused)
             for s, st in enumerate(strands)]
    out = ctx.classify(len(strands), n_refs, v_corRef, m_corRef, tax_rank, higher)
    return out, stats


def synthetic_lineage(n_refs, species_per_genus=4, genera_per_family=5):
    """A nested synthetic lineage for n_refs genomes: (rows for the lineage file, v_corRef at
    species rank, m_corRef [6][n_refs]).  Two consecutive genomes share a species."""
    g = np.arange(n_refs, dtype=np.uint32)
    species = 100000 + g // 2
    genus = 200000 + (g // 2) // species_per_genus
    family = 300000 + (g // 2) // (species_per_genus * genera_per_family)
    order = 400000 + family % 50
    clas = 500000 + order % 7
    phylum = 600000 + clas % 3
    m = np.stack([species, genus, family, order, clas, phylum]).astype(np.uint32)
    rows = [[f"ACC{i:06d}"] + [int(m[k, i]) for k in range(6)] for i in range(n_refs)]
    return rows, m[0].copy(), m
