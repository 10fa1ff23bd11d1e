"""Python access to the test oracle.  TEST INFRASTRUCTURE ONLY.

Two checkers live here:
  * `Oracle`     -- ctypes wrapper of oracle/liblime_oracle.so (the C restatement, lime_oracle.c);
  * `Reference`  -- runs the UNMODIFIED reference executables compiled into oracle/_ref/ by
                    oracle/Makefile on files in a scratch directory.
Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference leg may
import this module.  Nothing under lime_b200/ does.
"""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB = os.path.join(HERE, "liblime_oracle.so")
REF_DIR = os.path.join(HERE, "_ref")

CLUSTER_DT = np.dtype([("pStart", "<u8"), ("len", "<u8")])
PAIR_SIM_DT = np.dtype([("sim", "<f4"), ("idRef", "<u4")])


def build(ref=True):
    """make oracle (+ ref when /root/reference is present)."""
    subprocess.run(["make", "-s", "-C", HERE, "oracle"] + (["ref"] if ref else []), check=True)


def _p(a):
    return None if a is None else C.c_void_p(a.ctypes.data)


class Oracle:
    def __init__(self):
        if not os.path.exists(LIB):
            build(ref=False)
        L = C.CDLL(LIB)
        L.oracle_cluster_detect.restype = C.c_uint64
        L.oracle_cluster_detect.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64, C.c_uint32, C.c_uint32,
                                            C.c_void_p, C.c_uint64, C.POINTER(C.c_uint64)]
        L.oracle_similarity.restype = C.c_int
        L.oracle_similarity.argtypes = [C.c_void_p, C.c_uint64, C.c_void_p, C.c_void_p, C.c_uint32,
                                        C.c_uint32, C.c_void_p]
        L.oracle_choose.restype = C.c_uint64
        L.oracle_choose.argtypes = [C.c_void_p, C.c_uint32, C.c_uint32, C.c_uint32, C.c_float,
                                    C.c_void_p, C.c_void_p]
        L.oracle_fix_rank.restype = C.c_int64
        L.oracle_fix_rank.argtypes = [C.c_char_p, C.c_int, C.c_uint32, C.c_void_p, C.c_uint64, C.c_void_p]
        L.oracle_classify.restype = C.c_int
        L.oracle_classify.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint32,
                                      C.c_void_p, C.c_void_p, C.c_int, C.c_int,
                                      C.c_void_p, C.c_void_p, C.c_void_p]
        L.oracle_write_csv.restype = C.c_int
        L.oracle_write_csv.argtypes = [C.c_char_p, C.c_uint32, C.c_void_p, C.c_void_p, C.c_void_p]
        L.oracle_score_ebwt.restype = C.c_uint8
        L.oracle_score_ebwt.argtypes = [C.c_void_p, C.c_void_p]
        L.oracle_class_table.argtypes = [C.c_void_p]
        self.L = L

    def cluster_detect(self, lcp, da, n_reads, alpha):
        lcp = np.ascontiguousarray(lcp, dtype=np.uint32)
        da = np.ascontiguousarray(da, dtype=np.uint32)
        n = lcp.shape[0]
        mx = C.c_uint64()
        cnt = self.L.oracle_cluster_detect(_p(lcp), _p(da), n, n_reads, alpha, None, 0, C.byref(mx))
        out = np.zeros(cnt, dtype=CLUSTER_DT)
        self.L.oracle_cluster_detect(_p(lcp), _p(da), n, n_reads, alpha, _p(out), cnt, C.byref(mx))
        return out, int(mx.value)

    def similarity(self, clusters, da, ebwt, n_reads, n_refs):
        cl = np.ascontiguousarray(clusters, dtype=CLUSTER_DT)
        da = np.ascontiguousarray(da, dtype=np.uint32)
        if ebwt is not None:
            ebwt = np.ascontiguousarray(ebwt, dtype=np.uint8)
        sim = np.zeros((n_reads, n_refs), dtype=np.uint8)
        rc = self.L.oracle_similarity(_p(cl), cl.shape[0], _p(da), _p(ebwt), n_reads, n_refs, _p(sim))
        if rc != 0:
            raise RuntimeError(f"oracle_similarity rc={rc}")
        return sim

    def choose(self, sim, norm, beta):
        sim = np.ascontiguousarray(sim, dtype=np.uint8)
        n_reads, n_refs = sim.shape
        pos = np.zeros(n_reads, dtype=np.uint64)
        n = self.L.oracle_choose(_p(sim), n_reads, n_refs, norm, beta, None, _p(pos))
        bin_ = np.zeros(n, dtype=PAIR_SIM_DT)
        self.L.oracle_choose(_p(sim), n_reads, n_refs, norm, beta, _p(bin_), _p(pos))
        return bin_, pos

    def fix_rank(self, path, tax_rank, n_refs, higher=False):
        v = np.zeros(max(n_refs, 1), dtype=np.uint32)
        m = np.zeros((6, max(n_refs, 1)), dtype=np.uint32) if higher else None
        found = self.L.oracle_fix_rank(path.encode(), tax_rank, n_refs, _p(v), n_refs, _p(m))
        return (v[:n_refs] if found == n_refs else None), m, int(found)

    def classify(self, bins, poss, n_refs, v, m=None, tax_rank=1, higher=False):
        nf = len(bins)
        bins = [np.ascontiguousarray(b, dtype=PAIR_SIM_DT) for b in bins]
        poss = [np.ascontiguousarray(p, dtype=np.uint64) for p in poss]
        n_reads = poss[0].shape[0]
        bp = (C.c_void_p * nf)(*[b.ctypes.data for b in bins])
        pp = (C.c_void_p * nf)(*[p.ctypes.data for p in poss])
        v = np.ascontiguousarray(v, dtype=np.uint32)
        if m is not None:
            m = np.ascontiguousarray(m, dtype=np.uint32)
        typ = np.zeros(n_reads, dtype=np.uint8)
        tax = np.zeros(n_reads, dtype=np.uint32)
        sim = np.zeros(n_reads, dtype=np.float32)
        rc = self.L.oracle_classify(nf, bp, pp, n_reads, n_refs, _p(v), _p(m), tax_rank,
                                    1 if higher else 0, _p(typ), _p(tax), _p(sim))
        if rc != 0:
            raise RuntimeError(f"oracle_classify rc={rc}")
        return typ, tax, sim

    def write_csv(self, path, typ, tax, sim):
        typ = np.ascontiguousarray(typ, dtype=np.uint8)
        tax = np.ascontiguousarray(tax, dtype=np.uint32)
        sim = np.ascontiguousarray(sim, dtype=np.float32)
        assert self.L.oracle_write_csv(path.encode(), typ.shape[0], _p(typ), _p(tax), _p(sim)) == 0

    def score_ebwt(self, R, G):
        R = np.ascontiguousarray(R, dtype=np.uint8)
        G = np.ascontiguousarray(G, dtype=np.uint8)
        return int(self.L.oracle_score_ebwt(_p(R), _p(G)))


class Reference:
    """The unmodified reference executables (oracle/_ref), driven through files."""

    def __init__(self):
        self.dir = REF_DIR

    def available(self):
        return all(os.path.exists(os.path.join(self.dir, b)) for b in
                   ("ClusterLCP", "ClusterBWT_DA_ebwt1", "ClusterBWT_DA_ebwt0",
                    "Classify_higher0", "Classify_higher1"))

    def _run(self, exe, args, cwd):
        r = subprocess.run([os.path.join(self.dir, exe)] + [str(a) for a in args], cwd=cwd,
                           stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True)
        return r

    def cluster_lcp(self, fasta, n_reads, n_refs, alpha, threads=1):
        """-> (clusters sorted by pStart, aux dict)."""
        cwd = os.path.dirname(fasta)
        r = self._run("ClusterLCP", [fasta, n_reads, n_refs, alpha, threads], cwd)
        assert r.returncode == 0, r.stderr
        cl = np.fromfile(f"{fasta}.{alpha}.clrs", dtype=CLUSTER_DT)
        cl = cl[np.argsort(cl["pStart"], kind="stable")]
        stem = fasta[:fasta.find(".fasta")] if ".fasta" in fasta else fasta
        raw = open(stem + ".out", "rb").read()
        assert len(raw) == 28
        aux = {"numReads": int.from_bytes(raw[0:4], "little"), "numGenomes": int.from_bytes(raw[4:8], "little"),
               "alpha": int.from_bytes(raw[8:12], "little"), "maxLen": int.from_bytes(raw[12:20], "little"),
               "nClusters": int.from_bytes(raw[20:28], "little"), "raw": raw}
        return cl, aux

    def cluster_bwt_da(self, fasta, read_len, beta, threads=1, ebwt=True):
        cwd = os.path.dirname(fasta)
        r = self._run("ClusterBWT_DA_ebwt1" if ebwt else "ClusterBWT_DA_ebwt0", [fasta, read_len, beta, threads], cwd)
        if r.returncode != 0:
            return None, None, r
        return (np.fromfile(fasta + ".res.bin", dtype=PAIR_SIM_DT),
                np.fromfile(fasta + ".res.pos", dtype=np.uint64), r)

    def classify(self, stems, n_reads, n_refs, out_csv, lineage, tax_rank, threads=1, higher=False):
        cwd = os.path.dirname(out_csv)
        exe = "Classify_higher1" if higher else "Classify_higher0"
        r = self._run(exe, [len(stems)] + list(stems) + [n_reads, n_refs, out_csv, lineage, tax_rank, threads], cwd)
        return r
