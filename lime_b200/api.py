"""Python face of the C ABI (include/lime_cuda.h): one `LimeContext` per process and GPU.

This is plumbing for tests, the benchmark and the multi-GPU host; the product is the shared
library.  Arrays are numpy arrays (host) or raw device pointers (ints, e.g. tensor.data_ptr()).
"""
import ctypes as C
import numpy as np

from . import _lib
from ._lib import LimeError, MEM_HOST, MEM_DEVICE

CLUSTER_DT = np.dtype([("pStart", "<u8"), ("len", "<u8")])          # Tools.h:85-88
PAIR_SIM_DT = np.dtype([("sim", "<f4"), ("idRef", "<u4")])          # Tools.h:95-98


def _ptr(a):
    if a is None:
        return None
    if isinstance(a, int):
        return C.c_void_p(a)
    assert a.flags["C_CONTIGUOUS"]
    return C.c_void_p(a.ctypes.data)


class LimeContext:
    def __init__(self, device=0):
        self.lib = _lib.load()
        h = C.c_void_p()
        rc = self.lib.lime_init(C.byref(h), device)
        if rc != 0:
            raise LimeError(rc, (self.lib.lime_last_error(None) or b"").decode())
        self.h = h
        self._keep = []

    def close(self):
        if self.h:
            for p in getattr(self, "_pinned", {}).values():
                self.lib.lime_pinned_free(C.c_void_p(p[0]))
            self._pinned = {}
            self.lib.lime_destroy(self.h)
            self.h = None

    def pinned(self, key, shape, dtype):
        """A page-locked numpy array owned by the context (reused per key while large enough)."""
        self._pinned = getattr(self, "_pinned", {})
        dtype = np.dtype(dtype)
        nbytes = int(np.prod(shape)) * dtype.itemsize
        ent = self._pinned.get(key)
        if ent is None or ent[1] < nbytes:
            if ent is not None:
                self.lib.lime_pinned_free(C.c_void_p(ent[0]))
            ptr = self.lib.lime_pinned_alloc(max(nbytes, 1))
            if not ptr:
                raise MemoryError("lime_pinned_alloc failed")
            ent = (ptr, max(nbytes, 1))
            self._pinned[key] = ent
        buf = (C.c_uint8 * max(nbytes, 1)).from_address(ent[0])
        return np.frombuffer(buf, dtype=dtype, count=int(np.prod(shape))).reshape(shape)

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _ck(self, rc):
        if rc != 0:
            raise LimeError(rc, (self.lib.lime_last_error(self.h) or b"").decode())

    # ---- plumbing -----------------------------------------------------------------------
    def set_stream(self, stream_ptr):
        """Run this context's work on a caller stream.  A raw handle of 0 is CUDA's default stream
        (what torch.cuda.current_stream().cuda_stream returns unless a side stream is current): it is
        passed as cudaStreamLegacy (0x1) so that our kernels are ordered with the caller's work on it
        (events, NCCL collectives waited for on it).  None restores the context's own stream."""
        if stream_ptr is None:
            self._ck(self.lib.lime_set_stream(self.h, None))
        else:
            self._ck(self.lib.lime_set_stream(self.h, C.c_void_p(stream_ptr if stream_ptr else 1)))

    def launch_count(self):
        return int(self.lib.lime_launch_count(self.h))

    def prof_enable(self, on=True):
        self._ck(self.lib.lime_prof_enable(self.h, 1 if on else 0))

    def prof_reset(self):
        self._ck(self.lib.lime_prof_reset(self.h))

    def prof(self):
        out = {}
        for k in range(self.lib.lime_prof_kernels()):
            name, n, ms = C.c_char_p(), C.c_uint64(), C.c_double()
            self._ck(self.lib.lime_prof_get(self.h, k, C.byref(name), C.byref(n), C.byref(ms)))
            out[name.value.decode()] = {"launches": int(n.value), "ms": float(ms.value)}
        return out

    # ---- inputs -------------------------------------------------------------------------
    def set_alpha_limit(self, max_alpha):
        """lime_set_alpha_limit: promise alpha <= max_alpha for the strands loaded from now on; with
        1..255 host-mode loads ship lcp as min(lcp, 255) bytes (identical results).  0 = no promise."""
        self._ck(self.lib.lime_set_alpha_limit(self.h, int(max_alpha)))

    def set_alpha(self, alpha):
        """lime_set_alpha: the strands loaded from now on will be run with exactly this alpha; with 1..255
        host-mode loads ship one bit per position (lcp < alpha), identical results.  0 = off."""
        self._ck(self.lib.lime_set_alpha(self.h, int(alpha)))

    def load(self, lcp, da, ebwt, n_reads, n_refs, n_own=None, n_halo=0, pos_base=0, is_last=True,
             mem=MEM_HOST, n_total=None):
        """lime_load.  Host mode takes numpy arrays; device mode takes pointers + n_total."""
        if mem == MEM_HOST:
            lcp = np.ascontiguousarray(lcp, dtype=np.uint32)
            da = np.ascontiguousarray(da, dtype=np.uint32)
            if ebwt is not None:
                ebwt = np.ascontiguousarray(ebwt, dtype=np.uint8)
            n_total = lcp.shape[0]
            assert da.shape[0] == n_total and (ebwt is None or ebwt.shape[0] == n_total)
        if n_own is None:
            n_own = n_total - n_halo
        assert n_own + n_halo == n_total
        self._keep = [lcp, da, ebwt]
        self.n_reads, self.n_refs = int(n_reads), int(n_refs)
        self._ck(self.lib.lime_load(self.h, _ptr(lcp), _ptr(da), _ptr(ebwt), n_own, n_halo, pos_base,
                                    1 if is_last else 0, n_reads, n_refs, mem))

    def sim_bind(self, dev_ptr, rows, pitch):
        self._ck(self.lib.lime_sim_bind(self.h, C.c_void_p(dev_ptr) if dev_ptr else None, rows, pitch))

    def sim_pitch(self, n_refs):
        return int(self.lib.lime_sim_pitch(n_refs))

    # ---- step 1 ---------------------------------------------------------------------------
    def cluster_detect(self, alpha):
        n, mx = C.c_uint64(), C.c_uint64()
        self._ck(self.lib.lime_cluster_detect(self.h, alpha, C.byref(n), C.byref(mx)))
        self.n_clusters = int(n.value)
        return int(n.value), int(mx.value)

    def clusters(self):
        out = np.empty(self.n_clusters, dtype=CLUSTER_DT)
        self._ck(self.lib.lime_clusters_copy(self.h, _ptr(out), self.n_clusters))
        return out

    def clusters_set(self, cl, max_len=0):
        cl = np.ascontiguousarray(cl, dtype=CLUSTER_DT)
        self.n_clusters = cl.shape[0]
        self._ck(self.lib.lime_clusters_set(self.h, _ptr(cl), cl.shape[0], max_len))

    # ---- steps 1 + 2a fused ---------------------------------------------------------------
    def cluster_similarity(self, alpha, ebwt_mode, keep_clusters=False):
        n, mx = C.c_uint64(), C.c_uint64()
        self._ck(self.lib.lime_cluster_similarity(self.h, alpha, 1 if ebwt_mode else 0,
                                                  1 if keep_clusters else 0, C.byref(n), C.byref(mx)))
        self.n_clusters = int(n.value)
        return int(n.value), int(mx.value)

    # ---- step 2 ---------------------------------------------------------------------------
    def similarity(self, ebwt_mode):
        self._ck(self.lib.lime_similarity(self.h, 1 if ebwt_mode else 0))

    def sim_stats(self):
        out = (C.c_uint64 * 4)()
        self._ck(self.lib.lime_sim_stats(self.h, out))
        return {"warp_clusters": out[0], "block_clusters": out[1], "updates": out[2], "exact_pass": out[3]}

    def sim_matrix(self, rows=None):
        """Dense copy of the matrix; rows: the slab height in exchange mode (reads per rank)."""
        out = np.empty((self.n_reads if rows is None else rows, self.n_refs), dtype=np.uint8)
        self._ck(self.lib.lime_sim_copy(self.h, _ptr(out)))
        return out

    # ---- multi-GPU exchange -----------------------------------------------------------------
    def sim_export(self, read_begin, read_end, dev_ptr, cap):
        """Non-zero cells of rows [read_begin, read_end) as triples into device memory; returns the
        count (which may exceed cap: then nothing past cap was written and the caller retries)."""
        n = C.c_uint64()
        rc = self.lib.lime_sim_export(self.h, read_begin, read_end, C.c_void_p(dev_ptr) if dev_ptr else None, cap, C.byref(n))
        if rc != 0 and rc != _lib.ERR_NOMEM:
            self._ck(rc)
        return int(n.value)

    def sim_import(self, dev_ptr, n):
        self._ck(self.lib.lime_sim_import(self.h, C.c_void_p(dev_ptr) if dev_ptr else None, n))

    # ---- multi-GPU: rows on the read owner, updates routed as records -----------------------------
    def xchg_alloc(self, world, rank, reads_per_rank, n_refs, n_strands, flat_cap, inbox_cap):
        """Allocate this rank's slabs / update list / inbox; returns the 64-byte CUDA IPC handle to all-gather."""
        h = np.zeros(_lib.IPC_HANDLE_BYTES, dtype=np.uint8)
        self._ck(self.lib.lime_xchg_alloc(self.h, world, rank, reads_per_rank, n_refs, n_strands, flat_cap, inbox_cap, _ptr(h)))
        return h

    def xchg_open(self, handles=None):
        """Map every rank's inbox (handles: (world, 64) uint8); None = loopback on one GPU (tests)."""
        if handles is not None:
            handles = np.ascontiguousarray(handles, dtype=np.uint8).reshape(-1)
        self._ck(self.lib.lime_xchg_open(self.h, _ptr(handles)))

    def xchg_close(self):
        self._ck(self.lib.lime_xchg_close(self.h))

    def xchg_accumulate(self, alpha, ebwt_mode, strand):
        """Score the loaded shard into update records; routing to the owners is queued on the library's side stream."""
        self._ck(self.lib.lime_xchg_accumulate(self.h, alpha, 1 if ebwt_mode else 0, strand))

    def xchg_apply(self, strand):
        """Side stream: wait (on the device) for every source's records of this strand, add them to the slab."""
        self._ck(self.lib.lime_xchg_apply(self.h, strand))

    def xchg_choose(self, strand, norm, beta, n_own_reads):
        self._ck(self.lib.lime_xchg_choose(self.h, strand, norm, beta, n_own_reads))
        self._slot_shape = getattr(self, "_slot_shape", {})
        self._slot_shape[strand] = (None, n_own_reads)

    def xchg_status(self, strand):
        out = (C.c_uint64 * 8)()
        self._ck(self.lib.lime_xchg_status(self.h, strand, out))
        k = ("n_clusters", "max_len", "updates", "list_slots", "applied_t", "byte_sum", "n_records", "err")
        return {a: int(b) for a, b in zip(k, out)}

    # ---- one strand without a host synchronisation ------------------------------------------------
    def strand_run(self, slot, alpha, ebwt_mode, norm, beta):
        """lime_strand_run: lime_cluster_similarity + lime_choose of all reads, queued; checked at classify()."""
        self._ck(self.lib.lime_strand_run(self.h, slot, alpha, 1 if ebwt_mode else 0, norm, beta))
        self._slot_shape = getattr(self, "_slot_shape", {})
        self._slot_shape[slot] = (None, self.n_reads)

    def strand_status(self, slot):
        """-> (n_clusters, max_len, n_records, updates) of a slot that lime_classify has read back."""
        out = (C.c_uint64 * 4)()
        self._ck(self.lib.lime_strand_status(self.h, slot, out))
        if slot in getattr(self, "_slot_shape", {}):
            self._slot_shape[slot] = (int(out[2]), self._slot_shape[slot][1])
        return tuple(int(x) for x in out)

    def choose(self, slot, norm, beta, read_begin=0, read_end=None):
        if read_end is None:
            read_end = self.n_reads
        n = C.c_uint64()
        self._ck(self.lib.lime_choose(self.h, slot, norm, beta, read_begin, read_end, C.byref(n)))
        self._slot_shape = getattr(self, "_slot_shape", {})
        self._slot_shape[slot] = (int(n.value), read_end - read_begin)
        return int(n.value)

    def choose_result(self, slot):
        n_rec, n_reads = self._slot_shape[slot]
        if n_rec is None:
            n_rec = self.strand_status(slot)[2]
        bin_ = np.empty(n_rec, dtype=PAIR_SIM_DT)
        pos = np.empty(n_reads, dtype=np.uint64)
        self._ck(self.lib.lime_choose_copy(self.h, slot, _ptr(bin_), _ptr(pos)))
        return bin_, pos

    def slot_set(self, slot, bin_, pos):
        bin_ = np.ascontiguousarray(bin_, dtype=PAIR_SIM_DT)
        pos = np.ascontiguousarray(pos, dtype=np.uint64)
        self._ck(self.lib.lime_slot_set(self.h, slot, _ptr(bin_), bin_.shape[0], _ptr(pos), pos.shape[0]))
        self._slot_shape = getattr(self, "_slot_shape", {})
        self._slot_shape[slot] = (bin_.shape[0], pos.shape[0])

    # ---- step 3 ---------------------------------------------------------------------------
    def classify_on_device(self, n_files, n_refs, v_corRef, m_corRef=None, tax_rank=1, higher=False):
        """lime_classify without the copy to the host; fetch with classify_result_to_device()."""
        v = np.ascontiguousarray(v_corRef, dtype=np.uint32)
        m = None if m_corRef is None else np.ascontiguousarray(m_corRef, dtype=np.uint32)
        self._ck(self.lib.lime_classify(self.h, n_files, n_refs, _ptr(v), _ptr(m), tax_rank,
                                        1 if higher else 0, None, None, None))
        return self._slot_shape[0][1]

    def classify_result_to_device(self, type_ptr, tax_ptr, sim_ptr):
        """Device-to-device copy of the last classification (u8 / u32 / f32 arrays of the slot's reads)."""
        self._ck(self.lib.lime_classify_result(self.h, C.c_void_p(type_ptr), C.c_void_p(tax_ptr), C.c_void_p(sim_ptr), MEM_DEVICE))

    def classify(self, n_files, n_refs, v_corRef, m_corRef=None, tax_rank=1, higher=False):
        n_reads = self._slot_shape[0][1]
        v = np.ascontiguousarray(v_corRef, dtype=np.uint32)
        m = None if m_corRef is None else np.ascontiguousarray(m_corRef, dtype=np.uint32)
        # results land in page-locked arrays owned by the context: valid until the next classify()
        typ = self.pinned("cls_type", (n_reads,), np.uint8)
        tax = self.pinned("cls_tax", (n_reads,), np.uint32)
        sim = self.pinned("cls_sim", (n_reads,), np.float32)
        self._ck(self.lib.lime_classify(self.h, n_files, n_refs, _ptr(v), _ptr(m), tax_rank,
                                        1 if higher else 0, _ptr(typ), _ptr(tax), _ptr(sim)))
        return typ, tax, sim


# ---- host helpers (lime_host.h) -------------------------------------------------------------
def fix_rank(path, tax_rank, n_refs, higher=False):
    """FixRank (Classify.cpp:33-89) -> (v_corRef or None on size mismatch, m_corRef, found)."""
    lib = _lib.load()
    v = np.zeros(max(n_refs, 1), dtype=np.uint32)
    m = np.zeros((6, max(n_refs, 1)), dtype=np.uint32) if higher else None
    found = lib.lime_host_fix_rank(path.encode(), tax_rank, n_refs, _ptr(v), n_refs, _ptr(m))
    if found < 0:
        raise FileNotFoundError(path)
    return (v[:n_refs] if found == n_refs else None), (m[:, :n_refs].copy() if m is not None else None), int(found)


def read_gesa(path):
    """<fasta>.<numSeq>.gesa (egsa) -> (ebwt u8, lcp u32, da u32), EGSAtoBCR.cpp:72-91."""
    lib = _lib.load()
    n = lib.lime_host_read_gesa(path.encode(), None, None, None, 0)
    if n < 0:
        raise FileNotFoundError(path)
    ebwt, lcp, da = np.empty(n, np.uint8), np.empty(n, np.uint32), np.empty(n, np.uint32)
    lib.lime_host_read_gesa(path.encode(), _ptr(ebwt), _ptr(lcp), _ptr(da), n)
    return ebwt, lcp, da


def write_res_txt(path, bin_, pos):
    """BIN=0 text form of the .res.bin / .res.pos images (ClusterBWT_DA.cpp:414,424,440)."""
    lib = _lib.load()
    bin_ = np.ascontiguousarray(bin_, dtype=PAIR_SIM_DT)
    pos = np.ascontiguousarray(pos, dtype=np.uint64)
    rc = lib.lime_host_write_res_txt(path.encode(), _ptr(bin_), bin_.shape[0], _ptr(pos), pos.shape[0])
    if rc != 0:
        raise OSError(f"cannot write {path} ({rc})")


def read_res_txt(path, n_reads):
    """<fasta>.res.txt -> (.res.bin image, .res.pos image) holding the values of the text."""
    lib = _lib.load()
    pos = np.zeros(max(n_reads, 1), dtype=np.uint64)
    n = lib.lime_host_read_res_txt(path.encode(), n_reads, None, 0, _ptr(pos))
    if n < 0:
        raise OSError(f"cannot read {path} ({n})")
    bin_ = np.zeros(n, dtype=PAIR_SIM_DT)
    lib.lime_host_read_res_txt(path.encode(), n_reads, _ptr(bin_), n, _ptr(pos))
    return bin_, pos[:n_reads]


def write_csv(path, typ, tax, sim):
    lib = _lib.load()
    typ = np.ascontiguousarray(typ, dtype=np.uint8)
    tax = np.ascontiguousarray(tax, dtype=np.uint32)
    sim = np.ascontiguousarray(sim, dtype=np.float32)
    if lib.lime_host_write_csv(path.encode(), typ.shape[0], _ptr(typ), _ptr(tax), _ptr(sim)) != 0:
        raise OSError(f"cannot write {path}")
