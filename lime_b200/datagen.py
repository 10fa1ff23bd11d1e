"""Input synthesis in the reference file formats (.ebwt / .lcp / .da).

The upstream producers of these arrays (BCR, egsa, eGap) are network-fetched third-party tools
(Install_tools_preprocessing.sh:14-38) and are not available offline, so inputs are

  * `exact_collection`  -- tiny, exact: naive suffix sort of reads + genomes (correctness corpora);
  * `synth_arrays`      -- large, direct synthesis of the three arrays with the "runs model" of
                           SURVEY.md section 8(d) (benchmark shapes), deterministic per seed and identical
                           on CPU and CUDA (counter-based integer hashing, no float math).
"""
import numpy as np

# ------------------------------------------------------------------------------------------
# exact generator
# ------------------------------------------------------------------------------------------
_COMP = str.maketrans("ACGTRYSWKMBDHVN", "TGCAYRSWMKVHDBN")


def revcomp(s):
    """Reverse complement (what `seqtk seq -r` does in Preprocessing.sh:55,65)."""
    return s.translate(_COMP)[::-1]


def exact_collection(reads, genomes, terminator=b"$"):
    """eBWT / LCP / DA of the collection reads + genomes by naive suffix sorting.

    Document ids: reads 0..nR-1, genomes nR..nR+nG-1 (merge order of Preprocessing.sh:159-166).
    Every sequence has its own end marker, smaller than every letter and ordered by document id;
    lcp never extends over an end marker.  ebwt[i] is the symbol preceding suffix i, the
    terminator byte for a whole sequence.  Returns (ebwt u8, lcp u32, da u32).
    """
    seqs = list(reads) + list(genomes)
    suf = []
    for d, s in enumerate(seqs):
        for k in range(len(s) + 1):
            suf.append((s[k:], d, k))
    suf.sort(key=lambda t: (t[0], t[1]))
    n = len(suf)
    ebwt = np.empty(n, dtype=np.uint8)
    lcp = np.zeros(n, dtype=np.uint32)
    da = np.empty(n, dtype=np.uint32)
    term = terminator[0]
    prev = None
    for i, (s, d, k) in enumerate(suf):
        da[i] = d
        ebwt[i] = ord(seqs[d][k - 1]) if k > 0 else term
        if prev is not None:
            m = min(len(prev), len(s))
            j = 0
            while j < m and prev[j] == s[j]:
                j += 1
            lcp[i] = j
        prev = s
    return ebwt, lcp, da


def exact_collection_fast(reads, genomes, terminator=b"$"):
    """The same arrays as exact_collection for collections of 10^5..10^7 positions: prefix doubling and a
    vectorised LCP in numpy instead of sorting Python strings (tests/test_datagen.py checks the two against
    each other)."""
    seqs = list(reads) + list(genomes)
    D = len(seqs)
    lens = np.array([len(x) for x in seqs], dtype=np.int64)
    n = int(lens.sum()) + D
    text = np.empty(n, dtype=np.int64)                 # letters D + byte, end marker of document d = d
    doc = np.empty(n, dtype=np.uint32)
    off = np.empty(n, dtype=np.int64)
    starts = np.concatenate([[0], np.cumsum(lens + 1)[:-1]])
    for d, x in enumerate(seqs):
        a = int(starts[d]); L = len(x)
        text[a:a + L] = np.frombuffer(x.encode(), dtype=np.uint8).astype(np.int64) + D
        text[a + L] = d
        doc[a:a + L + 1] = d
        off[a:a + L + 1] = np.arange(L + 1)
    # prefix doubling
    rank = np.unique(text, return_inverse=True)[1].astype(np.int64)
    k = 1
    while True:
        nxt = np.full(n, -1, dtype=np.int64)
        nxt[:n - k] = rank[k:]
        order = np.lexsort((nxt, rank))
        ro, no = rank[order], nxt[order]
        new = np.empty(n, dtype=np.int64)
        new[order] = np.cumsum(np.concatenate([[0], ((ro[1:] != ro[:-1]) | (no[1:] != no[:-1])).astype(np.int64)]))
        rank = new
        if int(rank.max()) == n - 1:
            break
        k *= 2
    sa = np.empty(n, dtype=np.int64)
    sa[rank] = np.arange(n)
    # lcp of neighbours: compare column by column while the pair still matches (end markers are unique)
    lcp = np.zeros(n, dtype=np.uint32)
    act = np.arange(1, n)
    a, b = sa[act - 1], sa[act]
    depth = 0
    while act.size:
        ok = (np.maximum(a, b) + depth < n)
        ok[ok] = text[a[ok] + depth] == text[b[ok] + depth]
        ok &= text[np.minimum(a + depth, n - 1)] >= D            # never across an end marker
        lcp[act[ok]] += 1
        act, a, b = act[ok], a[ok], b[ok]
        depth += 1
    da = doc[sa]
    term = terminator[0]
    prev = np.where(off[sa] > 0, text[np.maximum(sa - 1, 0)] - D, term)
    return prev.astype(np.uint8), lcp, da.astype(np.uint32)


def random_dna(rng, n, alphabet="ACGT"):
    return "".join(rng.choice(list(alphabet), size=n))


def sample_reads(rng, genomes, n_reads, read_len, err=0.01, iupac=0.0):
    """Reads drawn from random genome positions with substitution errors / IUPAC symbols."""
    out, src = [], []
    for _ in range(n_reads):
        g = int(rng.integers(len(genomes)))
        G = genomes[g]
        p = int(rng.integers(0, max(1, len(G) - read_len + 1)))
        r = list(G[p:p + read_len])
        for j in range(len(r)):
            u = rng.random()
            if u < err:
                r[j] = "ACGT"[int(rng.integers(4))]
            elif u < err + iupac:
                r[j] = "RYSWKMBDHVN"[int(rng.integers(11))]
        out.append("".join(r))
        src.append(g)
    return out, src


# ------------------------------------------------------------------------------------------
# direct synthesis ("runs model")
# ------------------------------------------------------------------------------------------
_M64 = (1 << 64) - 1


def _mix(torch, x):
    """splitmix64 finaliser on int64 tensors (wrapping arithmetic, logical shifts emulated)."""
    def shr(v, k):
        return (v >> k) & ((1 << (64 - k)) - 1)
    x = x + (-7046029254386353131)                       # 0x9E3779B97F4A7C15
    x = (x ^ shr(x, 30)) * (-4658895280553007687)        # 0xBF58476D1CE4E5B9
    x = (x ^ shr(x, 27)) * (-7723592293110705685)        # 0x94D049BB133111EB
    return x ^ shr(x, 31)


def _u32(torch, idx, seed, stream):
    """uniform 32-bit integers (as int64 in [0, 2^32)) keyed by (seed, stream, idx)."""
    key = (seed * 0x9E3779B1 + stream * 0x85EBCA77) & 0x7FFFFFFFFFFFFFFF
    h = _mix(torch, idx ^ key)
    return (h >> 16) & 0xFFFFFFFF


def _geom_table(torch, p, kmax, device):
    """thresholds t[k] = floor(2^32 * (1 - (1-p)^(k+1))): Geometric(p) (failures before success)
    is searchsorted(t, u, right=True) for u uniform in [0, 2^32)."""
    q = 1.0 - p
    t = [int((1.0 - q ** (k + 1)) * 4294967296.0) for k in range(kmax)]
    return torch.tensor(t, dtype=torch.int64, device=device)


def synth_arrays(n, n_reads, n_refs, read_len, seed=20261104, alpha_min=12, p_long=1e-5,
                 trlcp=None, terminator=0x24, device="cpu", chunk=1 << 25, p_mixed=0.35,
                 max_seg=60000, lo=0, hi=None, read_frac=None):
    """Directly synthesise (ebwt u8, lcp i32, da i32) torch tensors of n positions on `device`.

    Segment lengths: P(L=1)=0.5, else 2+Geometric(1/6); with probability p_long uniform in
    [1024, max_seg].  lcp: segment head uniform in [0, alpha_min), interior alpha_min +
    Geometric(0.08) (so larger alphas split segments further), optionally clamped to trlcp.
    da: segment kind mixed / reads-only / genomes-only with a global read share of
    n_reads*(read_len+1)/n; a segment has a genome g; reads are drawn among reads with home
    genome g (r mod n_refs == g), genome members are g w.p. 0.9 else uniform.
    ebwt: segment-dominant base w.p. 0.9 else uniform ACGT, N w.p. 1e-4, other IUPAC w.p. 1e-6 on
    genome positions, terminator byte w.p. (n_reads+n_refs)/n.
    int32 views are bit-identical to the uint32 file format (all values < 2^31).
    read_frac: share of read positions (default n_reads*(read_len+1)/n); set it when the id range
    n_reads belongs to a larger job than these n positions (a shard of a multi-GPU run).
    lo / hi: return only positions [lo, hi) of the n-position model (every value is a function
    of (seed, position, segment), so a window equals the same slice of the full arrays).
    """
    import torch
    dev = torch.device(device)
    I64 = torch.int64
    f_r = min(0.95, n_reads * (read_len + 1) / float(n)) if read_frac is None else float(read_frac)
    q_m = min(0.8, max(0.2, f_r))
    x_r = f_r - p_mixed * q_m                      # reads-only share
    x_r = min(max(x_r, 0.0), 1.0 - p_mixed)
    T32 = 4294967296.0
    thr_mixed = int(p_mixed * T32)
    thr_ronly = int((p_mixed + x_r) * T32)
    thr_qm = int(q_m * T32)

    hi = n if hi is None else min(int(hi), n)
    lo = max(0, min(int(lo), hi))
    # ---- segments (only as many as cover [0, hi)) -------------------------------------------
    mean_len = 0.5 * 1 + 0.5 * 7.0 + p_long * (1024 + max_seg) / 2
    n_seg = int(hi / mean_len * 1.08) + 1024
    while True:
        sid = torch.arange(n_seg, dtype=I64, device=dev)
        u1 = _u32(torch, sid, seed, 1)
        u2 = _u32(torch, sid, seed, 2)
        u3 = _u32(torch, sid, seed, 3)
        geo6 = torch.searchsorted(_geom_table(torch, 1.0 / 6.0, 256, dev), u2, right=True)
        seg_len = torch.where(u1 < (1 << 31), torch.ones_like(sid), 2 + geo6)
        if p_long > 0:
            long_len = 1024 + (u2 % (max_seg - 1024 + 1))
            seg_len = torch.where(u3 < int(p_long * T32), long_len, seg_len)
        seg_end = torch.cumsum(seg_len, 0)
        if int(seg_end[-1]) >= hi:
            break
        n_seg *= 2
    k = int(torch.searchsorted(seg_end, torch.tensor([hi], dtype=I64, device=dev), right=False)[0]) + 1
    seg_end = seg_end[:k].clone()
    seg_end[-1] = min(int(seg_end[-1]), n)
    sid = sid[:k]
    seg_start = torch.cat([torch.zeros(1, dtype=I64, device=dev), seg_end[:-1]])
    uk = _u32(torch, sid, seed, 4)
    seg_kind = torch.where(uk < thr_mixed, 0, torch.where(uk < thr_ronly, 1, 2)).to(torch.int8)
    seg_g = (_u32(torch, sid, seed, 5) % n_refs).to(torch.int32)
    seg_base = (_u32(torch, sid, seed, 6) & 3).to(torch.int8)
    del u1, u2, u3, uk, geo6, seg_len

    ebwt = torch.empty(hi - lo, dtype=torch.uint8, device=dev)
    lcp = torch.empty(hi - lo, dtype=torch.int32, device=dev)
    da = torch.empty(hi - lo, dtype=torch.int32, device=dev)
    geo08 = _geom_table(torch, 0.08, 512, dev)
    acgt = torch.tensor([65, 67, 71, 84], dtype=torch.uint8, device=dev)
    iupac = torch.tensor([ord(c) for c in "RYSWKMBDHV"], dtype=torch.uint8, device=dev)
    thr_term = int(min(0.5, (f_r / (read_len + 1)) + n_refs / float(n)) * T32)
    reads_per_home = (n_reads + n_refs - 1) // n_refs

    for a in range(lo, hi, chunk):
        b = min(hi, a + chunk)
        pos = torch.arange(a, b, dtype=I64, device=dev)
        s = torch.searchsorted(seg_end, pos, right=True)
        head = seg_start[s] == pos
        # lcp
        ul = _u32(torch, pos, seed, 10)
        interior = alpha_min + torch.searchsorted(geo08, ul, right=True)
        v = torch.where(head, ul % max(alpha_min, 1), interior)
        if trlcp is not None:
            v = torch.clamp(v, max=trlcp)
        lcp[a - lo:b - lo] = v.to(torch.int32)
        # da
        kind = seg_kind[s]
        g = seg_g[s].to(I64)
        um = _u32(torch, pos, seed, 11)
        is_read = (kind == 1) | ((kind == 0) & (um < thr_qm))
        ur = _u32(torch, pos, seed, 12)
        r = g + n_refs * (ur % reads_per_home)
        r = torch.where(r < n_reads, r, r % n_reads)
        ug = _u32(torch, pos, seed, 13)
        gid = torch.where((ug & 0xFFFF) < 58982, g, (ug >> 16) % n_refs)     # 0.9 * 65536
        da[a - lo:b - lo] = torch.where(is_read, r, n_reads + gid).to(torch.int32)
        # ebwt
        ub = _u32(torch, pos, seed, 14)
        base = torch.where((ub & 0xFFFF) < 58982, seg_base[s].to(I64), (ub >> 16) & 3)
        sym = acgt[base]
        ui = _u32(torch, pos, seed, 15)
        sym = torch.where(ui < 429497, torch.full_like(sym, 78), sym)                       # 'N', 1e-4
        sym = torch.where((~is_read) & (ui >= 429497) & (ui < 429497 + 4295), iupac[(ui >> 3) % 10], sym)  # 1e-6
        ut = _u32(torch, pos, seed, 16)
        sym = torch.where(ut < thr_term, torch.full_like(sym, terminator), sym)
        ebwt[a - lo:b - lo] = sym
    if lo == 0 and hi > 0:
        lcp[0] = 0
    return ebwt, lcp, da


def write_strand(prefix, ebwt, lcp, da):
    """Write <prefix>.ebwt/.lcp/.da (little-endian u8/u32/u32, no header)."""
    def host(x):
        return x.cpu().numpy() if hasattr(x, "cpu") else np.asarray(x)
    if ebwt is not None:
        host(ebwt).astype(np.uint8, copy=False).tofile(prefix + ".ebwt")
    host(lcp).view(np.uint32).tofile(prefix + ".lcp")
    host(da).view(np.uint32).tofile(prefix + ".da")
