// k12_fused.cu -- steps 1 and 2a in ONE pass over lcp / da / ebwt:
// alpha-cluster detection (ClusterLCP.cpp:139-283) fused with the per-cluster similarity
// accumulation (clusterAnalyze, ClusterBWT_DA.cpp:256-358).  The cluster list (.clrs) is never
// materialised unless the caller asks for it (EMIT), and da / ebwt are read exactly once.
//
// A position i with lcp[i] < alpha is a "head"; a segment is [head, next head); a segment with at
// least one da < n_reads and one da >= n_reads is an alpha-cluster (k1 notes in DESIGN.md).  A CTA
// takes a tile of 4096 positions plus a 128-position look-ahead and owns the segments whose head
// is in the tile (the chunk ownership rule of ClusterLCP.cpp:150-202).
//
//   phase A  128-bit loads of lcp / da (32-bit of ebwt) for 4 positions per thread per round;
//            head bits (lcp < alpha) and read bits (da < n_reads) become two bitmaps in shared
//            memory; da / ebwt stay in registers.
//   phase B  one thread per 32-position word decides, with bit-parallel segmented OR-fills
//            (Kogge-Stone over the word, over the 32 words of a warp by ballot, over the 5 warps
//            through shared memory), which positions lie in a segment that holds both a read and
//            a genome: the "active" bitmap.  No per-cluster loop, no divergence.
//   phase C  active positions (about a fifth of the input) are compacted from registers into a
//            dense shared-memory stream of (id, class, cluster-start flag).
//   phase D  warps slide 32-lane windows over the compacted stream, one lane per position,
//            group equal ids with __match_any_sync, count symbol classes with ballots and add
//            t(read, genome) to the similarity matrix (same scoring as k2_similarity.cu).
//   Clusters longer than 32 positions or running past the look-ahead are rare; they are appended
//   to a list that the block-path kernel of k2_similarity.cu consumes afterwards.
//
// Algorithmic traffic: (8 + E) bytes per position, E = 1 with ebwt; + 16 B per cluster if EMIT.
#include "common.cuh"
#include "score.cuh"
#include <cstdlib>

namespace lime {
namespace k12 {

constexpr int THREADS = 256;
constexpr int TILE = 4096;
constexpr int HALO = 128;
constexpr int EXT = TILE + HALO;          // positions held per CTA
constexpr int WORDS = EXT / 32;           // 132
constexpr int TILE_WORDS = TILE / 32;     // 128
constexpr int NGRP = (WORDS + 31) / 32;   // 5 warps of word-owners
constexpr int GROUPS = EXT / 4;           // 1056 groups of 4 positions
constexpr int LOADS = (GROUPS + THREADS - 1) / THREADS;   // 5
constexpr int NONE = 1 << 30;
constexpr int MAXFAST = 32;               // longest cluster the window path takes
constexpr int QCAP = 128;                 // queued similarity updates per warp

// x_i |= x_{i-1} while c_i (position i continues the run of i-1); Kogge-Stone, 5 steps
__device__ __forceinline__ uint32_t fill_up(uint32_t x, uint32_t c)
{
    x |= (x << 1) & c;  c &= c << 1;
    x |= (x << 2) & c;  c &= c << 2;
    x |= (x << 4) & c;  c &= c << 4;
    x |= (x << 8) & c;  c &= c << 8;
    x |= (x << 16) & c;
    return x;
}
// x_i |= x_{i+1} while c_i (position i continues into i+1)
__device__ __forceinline__ uint32_t fill_down(uint32_t x, uint32_t c)
{
    x |= (x >> 1) & c;  c &= c >> 1;
    x |= (x >> 2) & c;  c &= c >> 2;
    x |= (x >> 4) & c;  c &= c >> 4;
    x |= (x >> 8) & c;  c &= c >> 8;
    x |= (x >> 16) & c;
    return x;
}
__device__ __forceinline__ uint32_t mask_lt(int k)   // bits [0, k), k may be <= 0 or >= 32
{
    return k <= 0 ? 0u : (k >= 32 ? 0xffffffffu : ((1u << k) - 1u));
}

struct Smem {
    uint32_t H[WORDS + 4];        // head bitmap (positions >= n_total are virtual heads)
    uint32_t R[WORDS + 4];        // read bitmap
    uint32_t A[WORDS + 4];        // active bitmap
    uint32_t base[WORDS + 4];     // exclusive prefix of popc(A)
    uint32_t cid[EXT];            // compacted document ids
    uint8_t  cmeta[EXT];          // compacted class | start << 4
    uint8_t  cls[256];            // symbol -> class
    uint2    q[THREADS / 32][QCAP]; // per-warp queue of updates {read, genome << 8 | t}
    uint32_t gHas[NGRP], gUpR[NGRP], gUpG[NGRP], gDnR[NGRP], gDnG[NGRP], gCnt[NGRP];
    uint32_t wsum[THREADS / 32];
    unsigned long long wmax[THREADS / 32];
    int E, hl;
    unsigned tile, tailR, tailG, slow_r, slow_g, smin, nact;
    unsigned cta_clusters, cta_updates;
    unsigned long long end, excl, cta_maxlen;
};

template <bool SCORE, bool EMIT, bool EBWT, bool EXACT32>
__global__ void __launch_bounds__(THREADS, 4) fused_kernel(K12Args P)
{
    __shared__ Smem sm;
    const int tid = threadIdx.x;
    const unsigned lane = lane_id(), warp = tid >> 5;
    const uint32_t lt_mask = (1u << lane) - 1u;

    if (tid == 0) {
        sm.tile = EMIT ? atomicAdd(P.tile_counter, 1u) : blockIdx.x;
        sm.E = NONE; sm.hl = -1; sm.tailR = 0; sm.tailG = 0;
        sm.cta_clusters = 0; sm.cta_updates = 0; sm.cta_maxlen = 0;
    }
    if (SCORE) {
        sm.cls[tid] = d_class_table.v[tid];
    }
    __syncthreads();
    const unsigned tile = sm.tile;
    const uint64_t a = (uint64_t)tile * TILE;                  // buffer-local start of the tile
    const int Lt = (int)min((uint64_t)TILE, P.n_own - a);      // owned positions in this tile
    const int Lext = (int)min((uint64_t)EXT, P.n_total - a);   // real positions in the window

    // ---- phase A: load, build the head / read bitmaps ------------------------------------------
    uint4 D[LOADS];
    uint32_t Eb[LOADS];
    {
        uint8_t *hB = reinterpret_cast<uint8_t *>(sm.H);
        uint8_t *rB = reinterpret_cast<uint8_t *>(sm.R);
#pragma unroll
        for (int k = 0; k < LOADS; k++) {
            const int q = tid + k * THREADS;
            if (q < GROUPS) {      // whole warps take this branch together (GROUPS % 32 == 0)
                const uint64_t p = a + 4ull * q;
                uint4 L;
                Eb[k] = 0;
                if (p + 4 <= P.n_total) {
                    L = ldcs128(P.lcp + p);
                    D[k] = ldcs128(P.da + p);
                    if (SCORE && EBWT) Eb[k] = __ldcs(reinterpret_cast<const uint32_t *>(P.ebwt + p));
                } else {
                    uint32_t l[4], d[4];
#pragma unroll
                    for (int e = 0; e < 4; e++) {
                        const bool ok = p + e < P.n_total;
                        l[e] = ok ? __ldg(P.lcp + p + e) : 0u;             // virtual head
                        d[e] = ok ? __ldg(P.da + p + e) : 0xffffffffu;     // neither read nor real
                        if (SCORE && EBWT && ok) Eb[k] |= (uint32_t)__ldg(P.ebwt + p + e) << (8 * e);
                    }
                    L = make_uint4(l[0], l[1], l[2], l[3]);
                    D[k] = make_uint4(d[0], d[1], d[2], d[3]);
                }
                const unsigned hn = (L.x < P.alpha) | ((L.y < P.alpha) << 1) | ((L.z < P.alpha) << 2) |
                                    ((L.w < P.alpha) << 3);
                const unsigned rn = (D[k].x < P.n_reads) | ((D[k].y < P.n_reads) << 1) |
                                    ((D[k].z < P.n_reads) << 2) | ((D[k].w < P.n_reads) << 3);
                const unsigned v = hn | (rn << 4);
                const unsigned o = __shfl_xor_sync(FULL, v, 1);
                if ((q & 1) == 0) hB[q >> 1] = (uint8_t)((v & 0xf) | ((o & 0xf) << 4));
                else              rB[q >> 1] = (uint8_t)((o >> 4) | (v & 0xf0));
            }
        }
    }
    __syncthreads();

    // ---- phase B1: per-word segment flags (threads 0..159 own one 32-position word each) ---------
    // A segment holds a read and a genome iff two neighbours inside it differ in kind:
    // X_i = (read_i != read_{i+1}) and i+1 is not a head.  One flag to spread over the segment.
    uint32_t H = 0xffffffffu, X = 0, V = 0, Rw = 0, Aw = 0;
    uint32_t lowmask = 0, himask = 0, upX = 0, dnX = 0;
    uint32_t HH = 0, YX = 0, ZX = 0;
    bool hasHead = true;
    const int w = tid;
    if (warp < NGRP) {
        if (w < WORDS) {
            V = mask_lt(Lext - w * 32);
            H = sm.H[w];
            Rw = sm.R[w];
            const uint32_t Hn = (H >> 1) | (sm.H[w + 1] << 31);
            const uint32_t Rn = (Rw >> 1) | (sm.R[w + 1] << 31);
            X = (Rw ^ Rn) & ~Hn & mask_lt(Lext - w * 32 - 1);
        }
        hasHead = H != 0;
        lowmask = hasHead ? ((H & (0u - H)) - 1u) : 0xffffffffu;          // below the first head
        himask = hasHead ? (0xffffffffu << (31 - __clz(H))) : 0xffffffffu; // at / above the last head
        upX = fill_up(X, ~H);
        dnX = fill_down(X, ~(H >> 1));
        HH = __ballot_sync(FULL, hasHead);
        const uint32_t TX = __ballot_sync(FULL, (X & himask) != 0);
        const uint32_t LX = __ballot_sync(FULL, (X & lowmask) != 0);
        YX = fill_up(TX, ~HH);
        ZX = fill_down(LX, ~HH);
        if (lane == 0) {
            sm.gHas[warp] = HH != 0;
            sm.gUpR[warp] = YX >> 31;
            sm.gDnR[warp] = ZX & 1u;
        }
        // first head at or after Lt (E), last head before Lt (hl)
        if (w < WORDS) {
            const uint32_t below_lt = mask_lt(Lt - w * 32);
            const uint32_t he = H & ~below_lt, hb = H & below_lt;
            if (he) atomicMin(&sm.E, w * 32 + __ffs(he) - 1);
            if (hb) atomicMax(&sm.hl, w * 32 + 31 - __clz(hb));
        }
    }
    __syncthreads();
    const int E = sm.E, hl = sm.hl;
    const bool open_tail = (E == NONE) && hl >= 0;
    const int owned_end = open_tail ? hl : (E == NONE ? 0 : E);
    if (warp < NGRP) {
        uint32_t cin = 0, cdn = 0;
        bool hbg = false;
        for (int k = 0; k < (int)warp; k++) {
            if (sm.gHas[k]) { cin = sm.gUpR[k]; hbg = true; }
            else            { cin |= sm.gUpR[k]; }
        }
        for (int k = NGRP - 1; k > (int)warp; k--) {
            if (sm.gHas[k]) cdn = sm.gDnR[k];
            else            cdn |= sm.gDnR[k];
        }
        const bool none_below = (HH & lt_mask) == 0;
        const bool none_above = lane == 31 || (HH >> (lane + 1)) == 0;
        const uint32_t ci = (((YX << 1) >> lane) & 1u) | (none_below ? cin : 0u);
        const uint32_t cd = (lane < 31 ? ((ZX >> (lane + 1)) & 1u) : 0u) | (none_above ? cdn : 0u);
        const bool head_before = !none_below || hbg;
        const uint32_t segX = upX | dnX | (ci ? lowmask : 0u) | (cd ? himask : 0u);
        const uint32_t PB = (hasHead ? ~lowmask : 0u) | (head_before ? lowmask : 0u);
        Aw = segX & PB & V & mask_lt(owned_end - w * 32);
        if (w < WORDS) {
            sm.A[w] = Aw;
            if (open_tail && (hl >> 5) == w) {          // in-window part of the segment that leaves the window
                const uint32_t mixed_in = (segX >> (hl & 31)) & 1u, kind = (Rw >> (hl & 31)) & 1u;
                sm.tailR = mixed_in | kind;
                sm.tailG = mixed_in | (kind ^ 1u);
            }
        }
        // a non-last shard whose last owned segment ends at the end of the buffer needs more halo
        if (w == 0 && !P.is_last && hl >= 0 && ((E != NONE && E == Lext && Lext < EXT))) atomicOr(P.err, 1u);
    }
    __syncthreads();

    // ---- phase B2 (rare): the last owned segment runs past the window; find its end ---------------
    // the whole CTA scans forward 1024 positions per step (ClusterLCP.cpp:246-264)
    bool tail_mixed = false;
    uint64_t tail_len = 0;
    if (open_tail) {
        uint64_t p0 = a + EXT;
        bool acc_r = false, acc_g = false;
        unsigned long long end = P.n_total;
        while (true) {
            if (p0 >= P.n_total) {
                if (!P.is_last && tid == 0) atomicOr(P.err, 1u);
                end = P.n_total;
                break;
            }
            if (tid == 0) sm.smin = 0xffffffffu;
            __syncthreads();
            uint32_t l[4], d[4];
            const uint64_t p = p0 + 4ull * tid;
            if (p + 4 <= P.n_total) {
                const uint4 lv = ldg128(P.lcp + p), dv = ldg128(P.da + p);
                l[0] = lv.x; l[1] = lv.y; l[2] = lv.z; l[3] = lv.w;
                d[0] = dv.x; d[1] = dv.y; d[2] = dv.z; d[3] = dv.w;
            } else {
#pragma unroll
                for (int e = 0; e < 4; e++) {
                    const bool ok = p + e < P.n_total;
                    l[e] = ok ? __ldg(P.lcp + p + e) : 0xffffffffu;
                    d[e] = ok ? __ldg(P.da + p + e) : 0u;
                }
            }
            unsigned hn = 0;
#pragma unroll
            for (int e = 0; e < 4; e++) hn |= (unsigned)(l[e] < P.alpha) << e;
            if (hn) atomicMin(&sm.smin, 4u * tid + __ffs(hn) - 1);
            __syncthreads();
            const unsigned m = sm.smin;
            bool r = false, g = false;
#pragma unroll
            for (int e = 0; e < 4; e++) {
                const bool in = (4u * tid + e < m) && (p + e < P.n_total);
                if (in) { if (d[e] < P.n_reads) r = true; else g = true; }
            }
            acc_r |= __syncthreads_or(r) != 0;
            acc_g |= __syncthreads_or(g) != 0;
            if (m != 0xffffffffu) { end = p0 + m; break; }
            p0 += 4ull * THREADS;
        }
        tail_mixed = (sm.tailR || acc_r) && (sm.tailG || acc_g);
        tail_len = end - (a + (uint64_t)hl);
    }

    // ---- phase B3: the clusters owned by this tile: count, longest, long list, records ------------
    // a thread walks the active heads of its word; the open tail (if mixed) belongs to word hl / 32
    const uint32_t heads = (warp < NGRP && w < TILE_WORDS) ? (H & Aw) : 0u;
    const bool own_tail = open_tail && tail_mixed && (hl >> 5) == w;
    unsigned count = __popc(heads) + (own_tail ? 1u : 0u);
    unsigned long long max_len = own_tail ? tail_len : 0ull;

    unsigned long long rec_base = 0;
    if (EMIT) {
        unsigned incl = count;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const unsigned n = __shfl_up_sync(FULL, incl, o);
            if (lane >= (unsigned)o) incl += n;
        }
        if (lane == 31) sm.wsum[warp] = incl;
        __syncthreads();
        unsigned warp_off = 0, tile_total = 0;
#pragma unroll
        for (int x = 0; x < THREADS / 32; x++) {
            const unsigned s = sm.wsum[x];
            if (x < (int)warp) warp_off += s;
            tile_total += s;
        }
        if (warp == 0) {
            const unsigned long long e = lookback_exclusive(P.tile_state, tile, tile_total);
            if (lane == 0) {
                sm.excl = e;
                if (tile == gridDim.x - 1) *P.n_clusters = e + tile_total;   // tickets: last tile id
            }
        }
        __syncthreads();
        rec_base = sm.excl + warp_off + (incl - count);
    }
    {
        // Only the last head of a word can start a cluster longer than 32 positions (any other has
        // a later head in the same word).  Without EMIT the lengths of the short clusters are not
        // needed here (phase D takes their maximum), so only that head is examined.
        uint32_t hb = heads;
        if (!EMIT && hb) {
            const uint32_t last = 0x80000000u >> __clz(H);
            hb = (hb & last) ? last : 0u;
        }
        while (hb) {
            const int b = __ffs(hb) - 1;
            hb &= hb - 1;
            const int i = w * 32 + b;
            uint32_t m = H & ~((2u << b) - 1u);          // next head in this word
            int j;
            if (m) j = w * 32 + __ffs(m) - 1;
            else {
                int x = w + 1;
                while ((m = sm.H[x]) == 0) x++;           // a head exists: i < owned_end <= E
                j = x * 32 + __ffs(m) - 1;
            }
            const int len = j - i;
            if (EMIT || len > MAXFAST) max_len = max(max_len, (unsigned long long)len);
            if (EMIT) {
                if (rec_base < P.out_cap)
                    *reinterpret_cast<ulonglong2 *>(&P.out[rec_base]) = make_ulonglong2(P.pos_base + a + i, (unsigned long long)len);
                rec_base++;
            }
            if (SCORE && len > MAXFAST) {
                // too long for the window path: hand over, drop its positions from the stream
                const unsigned long long slot = atomicAdd(P.long_count, 1ull);
                if (slot < P.long_cap)
                    *reinterpret_cast<ulonglong2 *>(&P.long_list[slot]) = make_ulonglong2(P.pos_base + a + i, (unsigned long long)len);
                for (int x = i >> 5; x <= (j - 1) >> 5; x++)
                    atomicAnd(&sm.A[x], ~(mask_lt(j - x * 32) & ~mask_lt(i - x * 32)));
            }
        }
        if (own_tail) {
            if (EMIT) {
                if (rec_base < P.out_cap)
                    *reinterpret_cast<ulonglong2 *>(&P.out[rec_base]) = make_ulonglong2(P.pos_base + a + hl, tail_len);
            }
            if (SCORE) {
                const unsigned long long slot = atomicAdd(P.long_count, 1ull);
                if (slot < P.long_cap)
                    *reinterpret_cast<ulonglong2 *>(&P.long_list[slot]) = make_ulonglong2(P.pos_base + a + hl, tail_len);
            }
        }
    }
    // tile totals: reduced in shared memory, ONE global atomic per CTA and scalar (same-address
    // atomics from every warp serialise in L2 and were measured to dominate the kernel)
    {
        unsigned c = count;
        unsigned long long mx = max_len;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            c += __shfl_xor_sync(FULL, c, o);
            mx = max(mx, __shfl_xor_sync(FULL, mx, o));
        }
        if (lane == 0 && warp < NGRP) {
            if (c) atomicAdd(&sm.cta_clusters, c);
            if (mx) atomicMax(&sm.cta_maxlen, mx);
        }
    }
    __syncthreads();
    if (tid == 0 && !(P.dbg & 1u)) {
        if (!EMIT && sm.cta_clusters) atomicAdd(P.n_clusters, (unsigned long long)sm.cta_clusters);
        if (!SCORE && sm.cta_maxlen) atomicMax(P.max_len, sm.cta_maxlen);
    }
    if (!SCORE) return;
    if (P.dbg & 8u) return;

    // ---- phase C: compact the active positions into a dense stream ----------------------------------
    {
        unsigned c = (warp < NGRP && w < WORDS) ? (unsigned)__popc(sm.A[w]) : 0u;
        unsigned incl = c;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const unsigned n = __shfl_up_sync(FULL, incl, o);
            if (lane >= (unsigned)o) incl += n;
        }
        if (warp < NGRP && lane == 31) sm.gCnt[warp] = incl;
        __syncthreads();
        if (warp < NGRP) {
            unsigned off = 0, total = 0;
#pragma unroll
            for (int k = 0; k < NGRP; k++) {
                const unsigned s = sm.gCnt[k];
                if (k < (int)warp) off += s;
                total += s;
            }
            if (w < WORDS) sm.base[w] = off + incl - c;
            if (tid == 0) sm.nact = total;
        }
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < LOADS; k++) {
        const int q = tid + k * THREADS;
        if (q < GROUPS) {
            const int wq = q >> 3, sh = (q & 7) * 4;
            const uint32_t aw = sm.A[wq];
            const uint32_t a4 = (aw >> sh) & 0xfu;
            if (a4) {
                uint32_t o = sm.base[wq] + __popc(aw & ((1u << sh) - 1u));
                const uint32_t h4 = (sm.H[wq] >> sh) & 0xfu;
                const uint32_t d[4] = {D[k].x, D[k].y, D[k].z, D[k].w};
#pragma unroll
                for (int e = 0; e < 4; e++) {
                    if ((a4 >> e) & 1u) {
                        const uint32_t st = (h4 >> e) & 1u;
                        const uint32_t cl = EBWT ? sm.cls[(Eb[k] >> (8 * e)) & 0xffu] : 0u;
                        sm.cid[o] = d[e];
                        sm.cmeta[o] = (uint8_t)(cl | (st << 4));
                        o++;
                    }
                }
            }
        }
    }
    __syncthreads();

    // ---- phase D: score the clusters of the stream, 32-lane windows --------------------------------
    const int n_act = (int)sm.nact;
    if (n_act == 0 || (P.dbg & 4u)) {            // CTA-uniform
        if (tid == 0 && sm.cta_maxlen && !(P.dbg & 1u)) atomicMax(P.max_len, sm.cta_maxlen);
        return;
    }
    SimAcc acc;
    acc.sim8 = P.sim8; acc.pitch8 = P.pitch8; acc.sim32 = P.sim32; acc.lo = P.lo; acc.hi = P.hi;
    acc.n_refs = P.n_refs; acc.carry = P.carry; acc.sum_t = nullptr;
    unsigned long long updates = 0;
    unsigned longest = 0;                        // longest cluster this warp scored (warp-uniform)

    // every warp takes a run of whole clusters: [first start >= warp*chunk, first start >= (warp+1)*chunk)
    auto first_start_ge = [&](int x) -> int {
        while (x < n_act) {
            const int k = x + (int)lane;
            const uint32_t m = __ballot_sync(FULL, k < n_act && (sm.cmeta[k] & 16u));
            if (m) return x + __ffs(m) - 1;
            x += 32;
        }
        return n_act;
    };
    // Updates are queued per warp and applied in bulk so that many atomics are in flight per lane
    // and their return values (needed for the carry check) are not waited for one by one.
    uint2 *q = sm.q[warp];
    int qn = 0;
    auto drain = [&]() {
        __syncwarp();
        if (P.dbg & 2u) { qn = 0; return; }
        for (int k = (int)lane; k < qn; k += 128) {
            uint2 e[4];
            bool on[4];
#pragma unroll
            for (int u = 0; u < 4; u++) { on[u] = k + 32 * u < qn; if (on[u]) e[u] = q[k + 32 * u]; }
            if (EXACT32) {
#pragma unroll
                for (int u = 0; u < 4; u++) if (on[u]) sim_add<true>(acc, e[u].x, e[u].y >> 8, e[u].y & 0xffu);
            } else {
                unsigned old[4];
                if (P.dbg & 16u) {                      // experiment: fire-and-forget (no carry check)
#pragma unroll
                    for (int u = 0; u < 4; u++) {
                        if (on[u]) {
                            const uint32_t g = e[u].y >> 8;
                            unsigned *wp = reinterpret_cast<unsigned *>(acc.sim8 + (size_t)e[u].x * acc.pitch8) + (g >> 2);
                            atomicAdd(wp, (e[u].y & 0xffu) << ((g & 3u) * 8u));
                        }
                    }
                    continue;
                }
#pragma unroll
                for (int u = 0; u < 4; u++) {
                    if (on[u]) {
                        const uint32_t g = e[u].y >> 8;
                        unsigned *wp = reinterpret_cast<unsigned *>(acc.sim8 + (size_t)e[u].x * acc.pitch8) + (g >> 2);
                        old[u] = atomicAdd(wp, (e[u].y & 0xffu) << ((g & 3u) * 8u));
                    }
                }
                bool carry = false;
#pragma unroll
                for (int u = 0; u < 4; u++)
                    if (on[u]) carry |= ((old[u] >> (((e[u].y >> 8) & 3u) * 8u)) & 0xffu) + (e[u].y & 0xffu) > 0xffu;
                if (carry) *acc.carry = 1u;
            }
        }
        updates += (unsigned long long)qn;
        qn = 0;
        __syncwarp();
    };
    const int chunk = (n_act + THREADS / 32 - 1) / (THREADS / 32);
    int i = first_start_ge((int)warp * chunk);
    const int end = first_start_ge((int)(warp + 1) * chunk);

    while (i < end) {
        const int idx = i + (int)lane;
        const bool valid = idx < end;
        const uint32_t id = valid ? sm.cid[idx] : 0xffffffffu;
        const uint32_t meta = valid ? sm.cmeta[idx] : 16u;                 // lanes past the end: lone starts
        const uint32_t sb = __ballot_sync(FULL, (meta & 16u) != 0);        // bit 0 is set: i is a start
        int take = min(32, end - i);
        if (end - i > 32 && !(sm.cmeta[i + 32] & 16u)) {
            const int ls = 31 - __clz(sb);                                // the last cluster is cut: leave it
            if (ls > 0) take = ls;
        }
        const bool act = (int)lane < take;
        const uint32_t le_mask = lt_mask | (1u << lane);
        const uint32_t start_bit = 0x80000000u >> __clz(sb & le_mask);     // first lane of my cluster
        const uint32_t later = sb & ~le_mask;
        const uint32_t upper = (later & (0u - later)) - 1u;                // lanes below the next start (all if none)
        const uint32_t cm = act ? (upper & ~(start_bit - 1u) & mask_lt(take)) : 0u;   // lanes of my cluster
        const uint32_t cls = meta & 15u;
        const bool is_read = id < P.n_reads;
        const unsigned mm = __match_any_sync(FULL, id) & cm;               // lanes with my id in my cluster
        const bool leader = act && (mm & lt_mask) == 0;
        longest = max(longest, __reduce_max_sync(FULL, (unsigned)__popc(cm)));

        // per-class counts of my id: popc(my group & lanes of that class)
        uint4 vec = make_uint4(0, 0, 0, 0);
        bool wide = false, iupac = false;
        if (!EBWT) {
            vec.x = (unsigned)__popc(mm);
        } else {
            const unsigned b0 = __ballot_sync(FULL, cls == 0u), b1 = __ballot_sync(FULL, cls == 1u);
            const unsigned b2 = __ballot_sync(FULL, cls == 2u), b3 = __ballot_sync(FULL, cls == 3u);
            const unsigned rest = __ballot_sync(FULL, act && cls >= 4u);
            vec.x = (unsigned)__popc(mm & b0) | ((unsigned)__popc(mm & b1) << 8) |
                    ((unsigned)__popc(mm & b2) << 16) | ((unsigned)__popc(mm & b3) << 24);
            wide = rest != 0;                                              // a class beyond A,C,G,T
            if (wide) {
                unsigned present = __reduce_or_sync(FULL, (act && cls >= 4u) ? (1u << cls) : 0u);
                iupac = (present & 0x7ff0u) != 0;
                while (present) {
                    const int c = __ffs(present) - 1;
                    present &= present - 1;
                    const unsigned b = __ballot_sync(FULL, act && cls == (uint32_t)c);
                    const unsigned v = (unsigned)__popc(mm & b) << ((c & 3) * 8);
                    if ((c >> 2) == 1) vec.y |= v; else if ((c >> 2) == 2) vec.z |= v; else vec.w |= v;
                }
            }
        }
        // reads wrap at 256 (:123,206), genomes saturate at 255 (:96-97,222-223): counts <= 32 here

        // every read leader against every genome leader of its cluster
        unsigned gl = __ballot_sync(FULL, leader && !is_read) & cm;
        if (!(leader && is_read)) gl = 0;
        while (__any_sync(FULL, gl != 0)) {
            const int src = __ffs(gl) - 1;                                 // -1 (lane 31) when idle: harmless
            const uint32_t Gx = __shfl_sync(FULL, vec.x, src);
            const uint32_t gid = __shfl_sync(FULL, id, src);
            uint32_t t;
            if (!EBWT) t = min(vec.x, Gx);
            else if (!wide) t = score_acgt(vec.x, Gx);
            else {
                uint4 Gv;
                Gv.x = Gx;
                Gv.y = __shfl_sync(FULL, vec.y, src); Gv.z = __shfl_sync(FULL, vec.z, src);
                Gv.w = __shfl_sync(FULL, vec.w, src);
                t = (iupac ? score_iupac(vec, Gv) : score_plain(vec, Gv)) & 0xffu;
            }
            if (gl == 0) t = 0;
            gl &= gl - 1;
            if (qn + 32 > QCAP) drain();
            const unsigned pm = __ballot_sync(FULL, t != 0);
            if (t) q[qn + __popc(pm & lt_mask)] = make_uint2(id, ((gid - P.n_reads) << 8) | t);
            qn += __popc(pm);
        }
        i += take;
    }
    drain();
    if (lane == 0) {
        if (updates) atomicAdd(&sm.cta_updates, (unsigned)updates);
        if (longest) atomicMax(&sm.cta_maxlen, (unsigned long long)longest);
    }
    __syncthreads();
    if (tid == 0 && !(P.dbg & 1u)) {
        if (sm.cta_updates) atomicAdd(P.n_updates, (unsigned long long)sm.cta_updates);
        if (sm.cta_maxlen) atomicMax(P.max_len, sm.cta_maxlen);
    }
}

} // namespace k12

static bool use_tile_kernel()
{
    static int v = -1;
    if (v < 0) { const char *e = getenv("LIME_K12"); v = (e && e[0] == 't') ? 1 : 0; }
    return v == 1;
}
// look-back descriptors / long-list slots: sized for the finer of the two decompositions
uint64_t k12_num_tiles(uint64_t n_own) { return k12s_num_chunks(n_own); }
uint64_t k12_long_cap(uint64_t n_total) { return n_total / (k12::MAXFAST + 1) + k12_num_tiles(n_total) + 16; }

cudaError_t launch_k12(const K12Args &a, bool score, bool emit, bool exact32, cudaStream_t s)
{
    if (!use_tile_kernel()) return launch_k12s(a, score, emit, exact32, s);
    const uint64_t tiles = (a.n_own + k12::TILE - 1) / k12::TILE;
    if (tiles == 0) return cudaSuccess;
    const unsigned grid = (unsigned)tiles;
    const bool ebwt = a.ebwt != nullptr;
#define LIME_K12(SC, EM, EB, EX) k12::fused_kernel<SC, EM, EB, EX><<<grid, k12::THREADS, 0, s>>>(a)
    if (!score)                        LIME_K12(false, true, false, false);
    else if (emit) {
        if (ebwt && !exact32)          LIME_K12(true, true, true, false);
        else if (ebwt && exact32)      LIME_K12(true, true, true, true);
        else if (!ebwt && !exact32)    LIME_K12(true, true, false, false);
        else                           LIME_K12(true, true, false, true);
    } else {
        if (ebwt && !exact32)          LIME_K12(true, false, true, false);
        else if (ebwt && exact32)      LIME_K12(true, false, true, true);
        else if (!ebwt && !exact32)    LIME_K12(true, false, false, false);
        else                           LIME_K12(true, false, false, true);
    }
#undef LIME_K12
    return cudaGetLastError();
}

} // namespace lime
