// k12_stream.cu -- steps 1 and 2a in ONE pass over lcp / da / ebwt, warp-autonomous version:
// alpha-cluster detection (ClusterLCP.cpp:139-283) fused with the per-cluster similarity
// accumulation (clusterAnalyze, ClusterBWT_DA.cpp:256-358).
//
// A position i with lcp[i] < alpha is a "head"; a segment is [head, next head); a segment with at
// least one da < n_reads and one da >= n_reads is an alpha-cluster.  The kernel is persistent: every
// WARP walks over chunks of 512 positions (16 words of 32) with a stride of 480, owns the segments
// whose head lies in the first 480 positions (the chunk ownership rule of ClusterLCP.cpp:150-202,
// the last word is look-ahead) and never synchronises with another warp:
//
//   stage    the next chunk of lcp / da / ebwt is copied global -> shared with cp.async while the
//            current chunk is scored (the warp's only DRAM reads: no load latency on the critical path).
//   phase A  lane = position inside a word: two ballots per word give the head word (lcp < alpha)
//            and the read word (da < n_reads); lane k keeps word k.
//   phase B  lane = word: bit-parallel segmented OR-fills (Kogge-Stone inside the word, the same
//            routine over the 32 words through ballots) decide which positions lie in a segment
//            that holds both a read and a genome: the "active" word.  No per-cluster loop.
//   phase C  active positions (about a fifth of the input) are compacted from the stage into the
//            warp's shared-memory stream (id, symbol class | cluster-start flag), word by word.
//   phase D  32-lane windows slide over the stream, one lane per position: equal ids are grouped
//            with __match_any_sync, symbol classes counted with ballots, every read leader meets
//            every genome leader of its cluster by shuffle; updates (read, genome, t) are fire-and-forget
//            adds to the similarity matrix (carries ruled out afterwards, see add_update).
//   Clusters longer than 32 positions or running past the look-ahead are rare; they are appended
//   to a list that the block-path kernel of k2_similarity.cu consumes afterwards.
//
// Algorithmic traffic: (8 + E) bytes per position, E = 1 with ebwt; + 16 B per cluster if EMIT.
#include "common.cuh"
#include "score.cuh"
#include <algorithm>

// compile-time experiment switches for the next round of tuning (value 1: row address per read leader,
// 2: arithmetic A/C/G/T classes, 4: packed 32-bit stream element, 8: phase C visits only the non-empty
// words); 0 in every shipped build -- the SASS
// of the default build is byte-identical with and without these blocks
#ifndef LIME_K12_EXP
#define LIME_K12_EXP 0
#endif

namespace lime {
namespace k12s {

constexpr int WARPS = 4;                  // per CTA; warps share nothing but the class table
constexpr int THREADS = WARPS * 32;
constexpr int WORDS = 16;                 // 32-position words per chunk
constexpr int CH = WORDS * 32;            // positions held per warp and chunk
constexpr int OWN = CH - 32;              // owned positions per chunk; one word of look-ahead is enough
                                          // for every cluster the window path takes (<= 32 positions)
constexpr int NONE = 1 << 30;
constexpr int MAXFAST = 32;               // longest cluster the window path takes
constexpr int SCAP = CH + 32;             // stream capacity incl. one window of padding

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

struct alignas(16) WarpSmem {
    uint32_t lcp[CH];             // stage of the next chunk, filled by cp.async while this one is scored
                                  // (dead after phase A: phase C keeps its per-word broadcast there)
    uint32_t da[CH];
    uint8_t  eb[CH];
    uint32_t cid[SCAP];           // compacted document ids
    uint16_t cmeta[SCAP];         // compacted symbol class | start << 8
};
struct Smem {
    WarpSmem w[WARPS];
    uint8_t  cls[256];            // symbol -> class
};

// 16-byte asynchronous global -> shared copy (LDGSTS), L1 bypassed
__device__ __forceinline__ void cp_async16(unsigned smem_addr, const void *gmem)
{
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_addr), "l"(gmem) : "memory");
}
// the same with an L2 eviction policy (createpolicy): the inputs are read once and should not push the
// similarity matrix out of L2
__device__ __forceinline__ void cp_async16_hint(unsigned smem_addr, const void *gmem, uint64_t pol)
{
    asm volatile("cp.async.cg.shared.global.L2::cache_hint [%0], [%1], 16, %2;" ::"r"(smem_addr), "l"(gmem), "l"(pol) : "memory");
}
// 0 evict_normal, 1 evict_first, 2 evict_last, 3 / 4 evict_last for one half / one quarter of the lines
__device__ __forceinline__ uint64_t l2_policy(int kind)
{
    uint64_t p;
    if (kind == 1) asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(p));
    else if (kind == 2) asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(p));
    else if (kind == 3) asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 0.5;" : "=l"(p));
    else if (kind == 4) asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 0.25;" : "=l"(p));
    else asm volatile("createpolicy.fractional.L2::evict_normal.b64 %0, 1.0;" : "=l"(p));
    return p;
}
__device__ __forceinline__ void red_add_hint(unsigned *addr, unsigned v, uint64_t pol)
{
    asm volatile("red.relaxed.gpu.global.add.L2::cache_hint.u32 [%0], %1, %2;" ::"l"(addr), "r"(v), "l"(pol) : "memory");
}
__device__ __forceinline__ void cp_async_wait_all()
{
    asm volatile("cp.async.wait_all;" ::: "memory");
}

// PEER: the rows of the similarity matrix live in their owners' slabs (one GPU per read range, mapped
// peer memory); every update is a fire-and-forget add that travels over NVLink / NVSwitch to the owner.
template <bool SCORE, bool EMIT, bool EBWT, bool EXACT32, bool PEER = false>
__global__ void __launch_bounds__(THREADS, 7) stream_kernel(const __grid_constant__ K12Args P)
{
    __shared__ Smem sm;
    const unsigned lane = lane_id(), warp = threadIdx.x >> 5;
    const uint32_t lt_mask = (1u << lane) - 1u;
    if (SCORE) {
        for (int i = threadIdx.x; i < 256; i += THREADS) sm.cls[i] = d_class_table.v[i];
    }
    __syncthreads();
    WarpSmem &ws = sm.w[warp];
    const int w = (int)lane;                                  // the word this lane owns in phase B

    const uint64_t n_chunks = (P.n_own + OWN - 1) / OWN;
    const uint64_t stride = (uint64_t)gridDim.x * WARPS;
    unsigned cl_count = 0;                                    // clusters counted by this lane
    unsigned long long max_len = 0;                           // longest cluster seen by this lane
    unsigned updates = 0;                                     // updates issued by this lane (a lane sees < 2^32 pairs)

    SimAcc acc;
    acc.sim8 = P.sim8; acc.pitch8 = P.pitch8; acc.sim32 = P.sim32; acc.lo = P.lo; acc.hi = P.hi;
    acc.n_refs = P.n_refs; acc.carry = P.carry; acc.sum_t = P.sum_t;
    acc.peer = PEER ? P.peer_dev : nullptr;
    // Matrix updates are fire-and-forget adds of t << 8*(g&3) to the 32-bit word that holds the uint8
    // cell: no return value to wait for.  A carry out of a byte (true cell sum > 255, where the
    // reference wraps, ClusterBWT_DA.cpp:183,247) lowers the byte sum of the matrix by 255 or 256, so
    // the host proves their absence afterwards by comparing the byte sum of the matrix with the
    // total of all t (sum_t) and otherwise reruns the pass into exact 32-bit cells.
    // L2 policies: the inputs are read once (evict_first), the lines of the similarity matrix should stay
    // (evict_last); together 4.23 -> 3.67 ms per step on the BASELINE workload.  Experiment switches:
    // dbg 1024 = inputs evict_normal, 2048 = adds without a policy, 4096 / 8192 = evict_last for 1/2, 1/4 of the lines
    const uint64_t pol_in = l2_policy((P.dbg & 1024u) ? 0 : 1);
    const uint64_t pol_mat = l2_policy((P.dbg & 4096u) ? 3 : (P.dbg & 8192u) ? 4 : 2);

    const bool red_mode = P.sum_t != nullptr;
    unsigned long long t_total = 0;                           // sum of t added by this lane
    auto add_update = [&](uint32_t r, uint32_t g, uint32_t t) {
        updates++;
        if (P.dbg & 2u) return;
        if (EXACT32) { sim_add<true>(acc, r, g, t); return; }
        if (PEER) {
            uint8_t *row = peer_row(P.peer, P.pitch8, r);
            if (P.dbg & 0x3f0u) {                   // experiments: where do the adds go
                const uint32_t q = r / P.peer.per;
                uint32_t rem = r - q * P.peer.per;
                uint32_t dst = q;
                if (P.dbg & 16u) dst = P.peer_self;                                  // all local
                if (P.dbg & 64u) dst = (P.peer_self + 1u) % P.peer_world;            // all to the next rank
                if (P.dbg & 128u) dst = (P.peer_self + 1u + q % 3u) % P.peer_world;  // three peers
                if (P.dbg & 256u) rem &= 0x3ffffu;                                   // 28 MB per owner
                row = P.peer.base[dst] + (size_t)rem * P.pitch8;
            }
            red_add_hint(reinterpret_cast<unsigned *>(row) + (g >> 2), t << ((g & 3u) * 8u), pol_mat);
            t_total += t;
            return;
        }
        if (red_mode) {
            unsigned *wp = reinterpret_cast<unsigned *>(acc.sim8 + (size_t)r * acc.pitch8) + (g >> 2);
            if (P.dbg & 2048u) atomicAdd(wp, t << ((g & 3u) * 8u));
            else red_add_hint(wp, t << ((g & 3u) * 8u), pol_mat);
            t_total += t;
            return;
        }
        acc.sum_t = nullptr;
        sim_add<false>(acc, r, g, t);
    };

    // Stage pipeline: the chunk after the current one is copied global -> shared asynchronously
    // (16 bytes per lane and copy) while the current one is scored; whole chunks only, the tail of
    // the buffer and unaligned buffers are loaded synchronously.
    const bool aligned = ((reinterpret_cast<uintptr_t>(P.lcp) | reinterpret_cast<uintptr_t>(P.da) |
                           (SCORE && EBWT ? reinterpret_cast<uintptr_t>(P.ebwt) : 0)) & 15u) == 0;
    // per-lane source / destination of its 16-byte pieces (piece k of lcp / da is 512 bytes further)
    const char *g_lcp = reinterpret_cast<const char *>(P.lcp) + 16 * lane;
    const char *g_da = reinterpret_cast<const char *>(P.da) + 16 * lane;
    const char *g_eb = reinterpret_cast<const char *>(P.ebwt) + 16 * lane;
    const unsigned s_lcp = (unsigned)__cvta_generic_to_shared(&ws.lcp[4 * lane]);
    const unsigned s_da = (unsigned)__cvta_generic_to_shared(&ws.da[4 * lane]);
    const unsigned s_eb = (unsigned)__cvta_generic_to_shared(&ws.eb[16 * lane]);
    auto stage_issue = [&](uint64_t cc) -> bool {
        if (cc >= n_chunks) return false;
        const uint64_t aa = cc * OWN;
        if (!aligned || aa + CH > P.n_total) return false;
        const char *pl = g_lcp + 4 * aa, *pd = g_da + 4 * aa;
#pragma unroll
        for (int k = 0; k < CH / 128; k++) {
            cp_async16_hint(s_lcp + 512 * k, pl + 512 * k, pol_in);
            cp_async16_hint(s_da + 512 * k, pd + 512 * k, pol_in);
        }
        if (SCORE && EBWT) {
            const char *pe = g_eb + aa;
#pragma unroll
            for (int k = 0; k < CH / 512; k++) cp_async16_hint(s_eb + 512 * k, pe + 512 * k, pol_in);
        }
        return true;
    };
    // masks of a whole chunk (every chunk but the last ones of the buffer)
    const uint32_t V_full = w < WORDS ? 0xffffffffu : 0u;
    const uint32_t XM_full = w < WORDS - 1 ? 0xffffffffu : (w == WORDS - 1 ? 0x7fffffffu : 0u);
    const uint32_t BL_full = mask_lt(OWN - w * 32);
    const uint64_t c_first = (uint64_t)blockIdx.x * WARPS + warp;
    bool staged = stage_issue(c_first);

    for (uint64_t c = c_first; c < n_chunks; c += stride) {
        const uint64_t a = c * OWN;                                  // buffer-local start of the chunk
        const int Lt = (int)min((uint64_t)OWN, P.n_own - a);         // owned positions
        const int Lext = (int)min((uint64_t)CH, P.n_total - a);      // real positions in the chunk

        // ---- phase A: head / read words (lane = position inside a word) ------------------------------
        uint32_t H = 0xffffffffu, R = 0;                             // lanes >= WORDS: beyond the chunk
        if (staged) {
            cp_async_wait_all();
            __syncwarp();
#pragma unroll
            for (int k = 0; k < WORDS; k++) {
                const uint32_t l = ws.lcp[32 * k + lane];
                const uint32_t d = ws.da[32 * k + lane];
                const uint32_t hk = __ballot_sync(FULL, l < P.alpha);
                const uint32_t rk = __ballot_sync(FULL, d < P.n_reads);
                if (w == k) { H = hk; R = rk; }
            }
        } else {
            // tail of the buffer (or unaligned buffers): guarded loads, da / ebwt put into the stage
            const uint32_t *pl = P.lcp + a + lane, *pd = P.da + a + lane;
            const uint8_t *pe = P.ebwt + a + lane;
#pragma unroll
            for (int k = 0; k < WORDS; k++) {
                const bool ok = 32 * k + (int)lane < Lext;
                const uint32_t l = ok ? __ldcs(pl + 32 * k) : 0u;              // virtual head
                const uint32_t d = ok ? __ldcs(pd + 32 * k) : 0xffffffffu;     // neither read nor real
                const uint32_t hk = __ballot_sync(FULL, l < P.alpha);
                const uint32_t rk = __ballot_sync(FULL, d < P.n_reads);
                if (w == k) { H = hk; R = rk; }
                if (SCORE) {
                    ws.da[32 * k + lane] = d;
                    if (EBWT) ws.eb[32 * k + lane] = ok ? __ldcs(pe + 32 * k) : (uint8_t)0;
                }
            }
            __syncwarp();
        }
        // the stage is free once da / ebwt of the active positions are compacted (phase C); without
        // scoring right away.  The copy lands while this chunk is scored.
        bool issued = false;
        auto issue_next = [&]() {
            if (!issued) { __syncwarp(); staged = stage_issue(c + stride); issued = true; }
        };
        if (!SCORE) issue_next();

        // ---- phase B: per-word segment flags (lane w owns word w) -------------------------------------
        // A segment holds a read and a genome iff two neighbours inside it differ in kind:
        // X_i = (read_i != read_{i+1}) and i+1 is not a head.  One flag to spread over the segment.
        const bool whole = Lext == CH && Lt == OWN;                       // warp-uniform
        const uint32_t V = whole ? V_full : mask_lt(Lext - w * 32);
        const uint32_t Hup = __shfl_down_sync(FULL, H, 1), Rup = __shfl_down_sync(FULL, R, 1);
        const uint32_t Hn = (H >> 1) | (Hup << 31), Rn = (R >> 1) | (Rup << 31);
        const uint32_t X = (R ^ Rn) & ~Hn & (whole ? XM_full : mask_lt(Lext - w * 32 - 1));   // the last position has no neighbour
        const bool hasHead = H != 0;
        const uint32_t lowmask = hasHead ? ((H & (0u - H)) - 1u) : 0xffffffffu;           // below the first head
        const uint32_t himask = hasHead ? (0xffffffffu << (31 - __clz(H))) : 0xffffffffu; // at / above the last head
        const uint32_t upX = fill_up(X, ~H);
        const uint32_t dnX = fill_down(X, ~(H >> 1));
        const uint32_t HH = __ballot_sync(FULL, hasHead);
        const uint32_t TX = __ballot_sync(FULL, (X & himask) != 0);
        const uint32_t LX = __ballot_sync(FULL, (X & lowmask) != 0);
        const uint32_t YX = fill_up(TX, ~HH);           // bit w: the run ending at the top of word w has X
        const uint32_t ZX = fill_down(LX, ~HH);         // bit w: the run starting at the bottom of word w has X
        const uint32_t ci = ((YX << 1) >> lane) & 1u;
        const uint32_t cd = lane < 31 ? ((ZX >> (lane + 1)) & 1u) : 0u;
        const bool head_before = (HH & lt_mask) != 0;
        const uint32_t segX = upX | dnX | (ci ? lowmask : 0u) | (cd ? himask : 0u);
        const uint32_t PB = (hasHead ? ~lowmask : 0u) | (head_before ? lowmask : 0u);
        // first head at or after Lt (E), last head before Lt (hl)
        int E = NONE, hl = -1;
        {
            const uint32_t below_lt = whole ? BL_full : mask_lt(Lt - w * 32);
            const uint32_t he = w < WORDS ? (H & ~below_lt) : 0u, hb = H & below_lt;   // lanes >= WORDS own no word
            const uint32_t Em = __ballot_sync(FULL, he != 0), Bm = __ballot_sync(FULL, hb != 0);
            if (Em) {
                const int src = __ffs(Em) - 1;
                E = src * 32 + __ffs(__shfl_sync(FULL, he, src)) - 1;
            }
            if (Bm) {
                const int src = 31 - __clz(Bm);
                hl = src * 32 + 31 - __clz(__shfl_sync(FULL, hb, src));
            }
        }
        const bool open_tail = (E == NONE) && hl >= 0;                    // warp-uniform
        const int owned_end = open_tail ? hl : (E == NONE ? 0 : E);
        uint32_t Aw = segX & PB & V & mask_lt(owned_end - w * 32);
        // a non-last shard whose last owned segment ends at the end of the buffer needs more halo
        if (lane == 0 && !P.is_last && hl >= 0 && E != NONE && E == Lext && Lext < CH) atomicOr(P.err, 1u);

        // ---- phase B2 (rare): the last owned segment runs past the chunk; find its end ----------------
        // the warp scans forward 128 positions per step (ClusterLCP.cpp:246-264)
        bool tail_mixed = false;
        unsigned long long tail_len = 0;
        if (open_tail) {
            const int src = hl >> 5;
            const uint32_t sx = __shfl_sync(FULL, segX, src), sr = __shfl_sync(FULL, R, src);
            const uint32_t mixed_in = (sx >> (hl & 31)) & 1u, kind = (sr >> (hl & 31)) & 1u;
            bool acc_r = (mixed_in | kind) != 0, acc_g = (mixed_in | (kind ^ 1u)) != 0;
            uint64_t p0 = a + CH;
            unsigned long long end = P.n_total;
            while (true) {
                if (p0 >= P.n_total) {
                    if (!P.is_last && lane == 0) atomicOr(P.err, 1u);
                    end = P.n_total;
                    break;
                }
                uint32_t l[4], d[4];
                const uint64_t p = p0 + 4ull * lane;
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
                const uint32_t hm = __ballot_sync(FULL, hn != 0);
                unsigned m = 0xffffffffu;                                // offset of the first head in this block
                if (hm) {
                    const int s2 = __ffs(hm) - 1;
                    m = 4u * s2 + __ffs(__shfl_sync(FULL, hn, s2)) - 1;
                }
                bool r = false, g = false;
#pragma unroll
                for (int e = 0; e < 4; e++) {
                    const bool in = (4u * lane + e < m) && (p + e < P.n_total);
                    if (in) { if (d[e] < P.n_reads) r = true; else g = true; }
                }
                acc_r |= __any_sync(FULL, r) != 0;
                acc_g |= __any_sync(FULL, g) != 0;
                if (hm) { end = p0 + m; break; }
                p0 += 128;
            }
            tail_mixed = acc_r && acc_g;
            tail_len = end - (a + (uint64_t)hl);
        }

        // ---- phase B3: the clusters owned by this chunk: count, longest, long list, records ------------
        // position of the first head in the words above this lane's word (every owned head has one)
        const uint32_t aboveW = lane < 31 ? (HH & (0xfffffffeu << lane)) : 0u;
        const int xw = aboveW ? __ffs(aboveW) - 1 : 31;
        const uint32_t Hx = __shfl_sync(FULL, H, xw);
        const int nh_above = aboveW ? xw * 32 + __ffs(Hx) - 1 : NONE;
        const uint32_t heads = H & Aw;                                    // heads of owned clusters
        const bool own_tail = open_tail && tail_mixed && (hl >> 5) == w;
        const unsigned count = __popc(heads) + (own_tail ? 1u : 0u);
        cl_count += count;
        if (own_tail) max_len = max(max_len, tail_len);

        if (!SCORE && !EMIT) {
            // counting pass of the two-pass record emission: clusters owned by this chunk
            const unsigned total = __reduce_add_sync(FULL, count);
            if (lane == 0) P.tile_state[c] = total;
            continue;
        }
        unsigned long long rec_base = 0;
        if (EMIT) {
            unsigned incl = count;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const unsigned n = __shfl_up_sync(FULL, incl, o);
                if (lane >= (unsigned)o) incl += n;
            }
            const unsigned total = __shfl_sync(FULL, incl, 31);
            const unsigned long long e0 = P.tile_state[c];                // exclusive prefix of the counting pass
            if (lane == 0 && c == n_chunks - 1) *P.n_clusters = e0 + total;
            rec_base = e0 + (incl - count);
            // every owned head: length to the next head, record in ascending pStart order
            uint32_t hb = heads;
            while (hb) {
                const int b = __ffs(hb) - 1;
                hb &= hb - 1;
                const uint32_t m = H & ~((2u << b) - 1u);              // next head in this word (b < 31 or m == 0)
                const int j = (b < 31 && m) ? w * 32 + __ffs(m) - 1 : nh_above;
                const int len = j - (w * 32 + b);
                max_len = max(max_len, (unsigned long long)len);
                if (rec_base < P.out_cap)
                    *reinterpret_cast<ulonglong2 *>(&P.out[rec_base]) = make_ulonglong2(P.pos_base + a + w * 32 + b, (unsigned long long)len);
                rec_base++;
            }
            if (own_tail && rec_base < P.out_cap)
                *reinterpret_cast<ulonglong2 *>(&P.out[rec_base]) = make_ulonglong2(P.pos_base + a + hl, tail_len);
        }
        if (SCORE) {
            // Only the last head of a word can start a cluster longer than 32 positions (any other has
            // a later head in the same word): hand those to the block path, drop them from the stream.
            bool cand = false;
            int ci0 = 0;
            if (hasHead) {
                const int b = 31 - __clz(H);
                ci0 = w * 32 + b;
                cand = ((heads >> b) & 1u) && (nh_above - ci0 > MAXFAST);
            }
            unsigned lm = __ballot_sync(FULL, cand);
            while (lm) {                                                  // rare
                const int src = __ffs(lm) - 1;
                lm &= lm - 1;
                const int i = __shfl_sync(FULL, ci0, src), j = __shfl_sync(FULL, nh_above, src);
                if (!EMIT && w == src) max_len = max(max_len, (unsigned long long)(j - i));
                if (lane == 0) {
                    const unsigned long long slot = atomicAdd(P.long_count, 1ull);
                    if (slot < P.long_cap)
                        *reinterpret_cast<ulonglong2 *>(&P.long_list[slot]) = make_ulonglong2(P.pos_base + a + i, (unsigned long long)(j - i));
                }
                Aw &= ~(mask_lt(j - w * 32) & ~mask_lt(i - w * 32));
            }
            if (open_tail && tail_mixed && lane == 0) {
                const unsigned long long slot = atomicAdd(P.long_count, 1ull);
                if (slot < P.long_cap)
                    *reinterpret_cast<ulonglong2 *>(&P.long_list[slot]) = make_ulonglong2(P.pos_base + a + hl, tail_len);
            }
        }
        if (!SCORE) continue;
        if (P.dbg & 8u) { issue_next(); continue; }

        // ---- phase C: compact the active positions into the warp's stream ------------------------------
        // lane = position inside a word again: per word one broadcast of (active, head, offset), the
        // active lanes copy id / symbol / start flag from the stage; words without active positions
        // are skipped (warp-uniform)
        int n_act;
        {
            const unsigned cnt = (unsigned)__popc(Aw);
            unsigned incl = cnt;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const unsigned n = __shfl_up_sync(FULL, incl, o);
                if (lane >= (unsigned)o) incl += n;
            }
            n_act = (int)__shfl_sync(FULL, incl, 31);
            uint4 *info = reinterpret_cast<uint4 *>(ws.lcp);      // per word: active, head, exclusive prefix, -
            if (w < WORDS) info[w] = make_uint4(Aw, H, incl - cnt, 0u);
            __syncwarp();
            if (n_act) {
#if LIME_K12_EXP & 8
                // experiment (not built by default): visit only the words that hold active positions
                for (unsigned nzw = __ballot_sync(FULL, w < WORDS && Aw != 0u); nzw; nzw &= nzw - 1u) {
                    const int k = __ffs(nzw) - 1;
                    const uint4 inf = info[k];
#else
#pragma unroll
                for (int k = 0; k < WORDS; k++) {
                    const uint4 inf = info[k];
                    if (inf.x == 0u) continue;
#endif
                    if ((inf.x >> lane) & 1u) {
                        const uint32_t o = inf.z + __popc(inf.x & lt_mask);
#if !(LIME_K12_EXP & 4)
                        ws.cid[o] = ws.da[32 * k + lane];
#endif
#if LIME_K12_EXP & 2
                        // experiment (not built by default): A,C,G,T mapped arithmetically, the table only for the rest
                        uint32_t cl0 = 0u;
                        if (EBWT) {
                            const uint32_t sy = ws.eb[32 * k + lane], d = sy - 0x41u;
                            const uint32_t x = (sy >> 1) & 3u;
                            cl0 = (d < 20u && ((0x80045u >> d) & 1u)) ? (x ^ (x >> 1)) : (uint32_t)sm.cls[sy];
                        }
#else
                        const uint32_t cl0 = EBWT ? (uint32_t)sm.cls[ws.eb[32 * k + lane]] : 0u;
#endif
#if LIME_K12_EXP & 4
                        // experiment (not built by default): one packed element id | class << 27 | start << 31
                        // (ids < 2^27 - 1, checked by the launcher): one store here, one load in phase D
                        ws.cid[o] = ws.da[32 * k + lane] | (cl0 << 27) | (((inf.y >> lane) & 1u) << 31);
#else
                        ws.cmeta[o] = (uint16_t)(cl0 | (((inf.y >> lane) & 1u) << 8));
#endif
                    }
                }
            }
        }
        issue_next();                                                     // includes the __syncwarp
        if (n_act == 0) continue;                                         // warp-uniform
        ws.cid[n_act + lane] = 0xffffffffu;                               // one window of padding: lone starts
#if !(LIME_K12_EXP & 4)
        ws.cmeta[n_act + lane] = 0x100u;
#endif
        __syncwarp();
        if (P.dbg & 4u) continue;

        // ---- phase D: score the clusters of the stream, 32-lane windows --------------------------------
        unsigned longest = 0;                                             // warp-uniform
        for (int i = 0; i < n_act;) {
#if LIME_K12_EXP & 4
            const uint32_t pk = ws.cid[i + lane];
            uint32_t id = pk & 0x07ffffffu;
            if (id == 0x07ffffffu) id = 0xffffffffu;                      // padding
            const uint32_t meta = ((pk >> 27) & 15u) | ((pk >> 31) << 8);
#define K12_NEXT_IS_START ((ws.cid[i + 32] >> 31) != 0u)
#else
            uint32_t id = ws.cid[i + lane];
            const uint32_t meta = ws.cmeta[i + lane];
#define K12_NEXT_IS_START (ws.cmeta[i + 32] & 0x100u)
#endif
            uint32_t sb = __ballot_sync(FULL, (meta & 0x100u) != 0);      // bit 0 is set: i is a start
            if (!K12_NEXT_IS_START) {
                // the last cluster of the window is cut: leave it to the next window (turn its lanes
                // into lone starts; a cluster has at most 32 positions, so it does not start at lane 0)
                const int ls = 31 - __clz(sb);
                sb |= 0xffffffffu << ls;
                if ((int)lane >= ls) id = 0xffffffffu;
                i += ls;
            } else {
                i += 32;
            }
            const uint32_t le_mask = lt_mask | (1u << lane);
            const uint32_t start_bit = 0x80000000u >> __clz(sb & le_mask);     // first lane of my cluster
            const uint32_t later = sb & ~le_mask;
            const uint32_t upper = (later & (0u - later)) - 1u;                // lanes below the next start (all if none)
            const uint32_t cm = upper & ~(start_bit - 1u);                     // lanes of my cluster
            const bool is_read = id < P.n_reads;
            const unsigned mm = __match_any_sync(FULL, id) & cm;               // lanes with my id in my cluster
            const bool leader = (mm & lt_mask) == 0;
            longest = max(longest, __reduce_max_sync(FULL, (unsigned)__popc(cm)));

            // per-class counts of my id: popc(my group & lanes of that class)
            uint4 vec = make_uint4(0, 0, 0, 0);
            bool wide = false, iupac = false;
            if (!EBWT) {
                vec.x = (unsigned)__popc(mm);
            } else {
                const uint32_t cls = meta & 0xffu;
                const unsigned c0 = __ballot_sync(FULL, (cls & 1u) != 0), c1 = __ballot_sync(FULL, (cls & 2u) != 0);
                const unsigned rest = __ballot_sync(FULL, cls >= 4u && id != 0xffffffffu);
                vec.x = (unsigned)__popc(mm & ~c1 & ~c0) | ((unsigned)__popc(mm & ~c1 & c0) << 8) |
                        ((unsigned)__popc(mm & c1 & ~c0) << 16) | ((unsigned)__popc(mm & c1 & c0) << 24);
                wide = rest != 0;                                              // a class beyond A,C,G,T
                if (wide) {
                    // exact per-class counts for every class present in the window
                    const bool real = id != 0xffffffffu;
                    unsigned present = __reduce_or_sync(FULL, real ? (1u << cls) : 0u);
                    iupac = (present & 0x7ff0u) != 0;
                    vec = make_uint4(0, 0, 0, 0);
                    while (present) {
                        const int cc = __ffs(present) - 1;
                        present &= present - 1;
                        const unsigned b = __ballot_sync(FULL, real && cls == (uint32_t)cc);
                        const unsigned v = (unsigned)__popc(mm & b) << ((cc & 3) * 8);
                        if ((cc >> 2) == 0) vec.x |= v; else if ((cc >> 2) == 1) vec.y |= v;
                        else if ((cc >> 2) == 2) vec.z |= v; else vec.w |= v;
                    }
                }
            }
            // reads wrap at 256 (:123,206), genomes saturate at 255 (:96-97,222-223): counts <= 32 here

            // every read leader against every genome leader of its cluster
            unsigned gl = __ballot_sync(FULL, leader && !is_read) & cm;
            if (!(leader && is_read)) gl = 0;
#if LIME_K12_EXP & 1
            // experiment (not built by default): the row address once per read leader instead of once per pair
            unsigned *row_w = reinterpret_cast<unsigned *>(acc.sim8 + (size_t)(gl ? id : 0u) * acc.pitch8);
#endif
            while (__any_sync(FULL, gl != 0)) {
                const int src = gl ? 31 - __clz(gl) : 0;
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
                else gl ^= 1u << src;
#if LIME_K12_EXP & 1
                if (t) {
                    if (!PEER && !EXACT32 && red_mode && !(P.dbg & 0x802u)) {
                        const uint32_t g = gid - P.n_reads;
                        updates++;
                        red_add_hint(row_w + (g >> 2), t << ((g & 3u) * 8u), pol_mat);
                        t_total += t;
                    } else add_update(id, gid - P.n_reads, t);
                }
#else
                if (t) add_update(id, gid - P.n_reads, t);
#endif
            }
        }
        max_len = max(max_len, (unsigned long long)longest);
    }
    // ---- totals of this warp: one atomic per warp and scalar -------------------------------------------
    {
        unsigned c2 = cl_count;
        unsigned long long mx = max_len, up = updates, tt = t_total;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            c2 += __shfl_xor_sync(FULL, c2, o);
            mx = max(mx, __shfl_xor_sync(FULL, mx, o));
            up += __shfl_xor_sync(FULL, up, o);
            tt += __shfl_xor_sync(FULL, tt, o);
        }
        if (lane == 0 && !(P.dbg & 1u)) {
            if (!EMIT && c2) atomicAdd(P.n_clusters, (unsigned long long)c2);
            if (mx) atomicMax(P.max_len, mx);
            if (SCORE && up) atomicAdd(P.n_updates, up);
            if (SCORE && tt) atomicAdd(P.sum_t, tt);
        }
    }
}

// ---- exclusive prefix over the per-chunk cluster counts (between the counting and the emitting pass) ----
constexpr int SCAN_THREADS = 1024;
__global__ void __launch_bounds__(SCAN_THREADS) chunk_scan_kernel(unsigned long long *a, uint64_t n, unsigned long long *total)
{
    __shared__ unsigned long long part[SCAN_THREADS];
    const uint64_t per = (n + SCAN_THREADS - 1) / SCAN_THREADS;
    const uint64_t lo = min(n, per * threadIdx.x), hi = min(n, lo + per);
    unsigned long long sum = 0;
    for (uint64_t i = lo; i < hi; i++) sum += a[i];
    part[threadIdx.x] = sum;
    __syncthreads();
    for (int o = 1; o < SCAN_THREADS; o <<= 1) {                 // Hillis-Steele over the 1024 partial sums
        const unsigned long long v = threadIdx.x >= (unsigned)o ? part[threadIdx.x - o] : 0ull;
        __syncthreads();
        part[threadIdx.x] += v;
        __syncthreads();
    }
    unsigned long long run = part[threadIdx.x] - sum;
    for (uint64_t i = lo; i < hi; i++) { const unsigned long long v = a[i]; a[i] = run; run += v; }
    if (threadIdx.x == SCAN_THREADS - 1 && total) *total = part[threadIdx.x];
}

// ---- byte sum of the matrix (carry check of the fire-and-forget adds) -------------------------------
constexpr int SUM_THREADS = 256;
__global__ void __launch_bounds__(SUM_THREADS) bytesum_kernel(const uint4 *p, uint64_t n16, unsigned long long *out)
{
    unsigned acc4 = 0;                                          // <= 2^32 / 1020 vectors per thread
    unsigned long long total = 0;
    for (uint64_t i = (uint64_t)blockIdx.x * SUM_THREADS + threadIdx.x; i < n16; i += (uint64_t)gridDim.x * SUM_THREADS) {
        const uint4 v = __ldcs(p + i);
        acc4 = __dp4a(v.x, 0x01010101u, acc4); acc4 = __dp4a(v.y, 0x01010101u, acc4);
        acc4 = __dp4a(v.z, 0x01010101u, acc4); acc4 = __dp4a(v.w, 0x01010101u, acc4);
        if (acc4 > 0xf0000000u) { total += acc4; acc4 = 0; }
    }
    total += acc4;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) total += __shfl_xor_sync(FULL, total, o);
    __shared__ unsigned long long part[SUM_THREADS / 32];
    if ((threadIdx.x & 31) == 0) part[threadIdx.x >> 5] = total;
    __syncthreads();
    if (threadIdx.x == 0) {
        unsigned long long t = 0;
        for (int k = 0; k < SUM_THREADS / 32; k++) t += part[k];
        if (t) atomicAdd(out, t);
    }
}

} // namespace k12s

// bytes must be a multiple of 16 and p 16-byte aligned (the matrix pitch is)
cudaError_t launch_bytesum(const uint8_t *p, uint64_t bytes, unsigned long long *out, cudaStream_t s)
{
    const uint64_t n16 = bytes / 16;
    if (!n16) return cudaSuccess;
    const unsigned grid = (unsigned)std::min<uint64_t>((n16 + k12s::SUM_THREADS - 1) / k12s::SUM_THREADS, 148ull * 16);
    k12s::bytesum_kernel<<<grid, k12s::SUM_THREADS, 0, s>>>(reinterpret_cast<const uint4 *>(p), n16, out);
    return cudaGetLastError();
}

uint64_t k12s_num_chunks(uint64_t n_own) { return (n_own + k12s::OWN - 1) / k12s::OWN; }

template <bool SC, bool EM, bool EB, bool EX, bool PE = false>
static unsigned grid_of(uint64_t n_own)
{
    static int ctas_per_sm = 0, sms = 0;
    if (!ctas_per_sm) {
        int dev = 0;
        if (cudaGetDevice(&dev) != cudaSuccess) return 0;
        if (cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess) return 0;
        if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&ctas_per_sm, k12s::stream_kernel<SC, EM, EB, EX, PE>, k12s::THREADS, 0) != cudaSuccess) return 0;
        if (ctas_per_sm < 1) ctas_per_sm = 1;
    }
    const uint64_t chunks = k12s_num_chunks(n_own);
    const uint64_t want = (chunks + k12s::WARPS - 1) / k12s::WARPS;
    return (unsigned)std::min<uint64_t>(want, (uint64_t)sms * ctas_per_sm);
}

template <bool SC, bool EM, bool EB, bool EX, bool PE = false>
static cudaError_t launch_variant(const K12Args &a, cudaStream_t s)
{
    // persistent grid, warps stride over chunks
    const unsigned grid = grid_of<SC, EM, EB, EX, PE>(a.n_own);
    if (!grid) return cudaErrorUnknown;
    k12s::stream_kernel<SC, EM, EB, EX, PE><<<grid, k12s::THREADS, 0, s>>>(a);
    return cudaGetLastError();
}

#define LIME_K12S_DISPATCH(CALL)                                                              \
    if (!score)                        return emit ? CALL(false, true, false, false) : CALL(false, false, false, false); \
    if (emit) {                                                                               \
        if (ebwt && !exact32)          return CALL(true, true, true, false);                 \
        if (ebwt && exact32)           return CALL(true, true, true, true);                  \
        if (!ebwt && !exact32)         return CALL(true, true, false, false);                \
        return CALL(true, true, false, true);                                                 \
    }                                                                                         \
    if (ebwt && !exact32)              return CALL(true, false, true, false);                \
    if (ebwt && exact32)               return CALL(true, false, true, true);                 \
    if (!ebwt && !exact32)             return CALL(true, false, false, false);               \
    return CALL(true, false, false, true);

// per-chunk cluster counts / record offsets (EMIT) are per chunk; the long-cluster list holds at most one record per
// 33 positions plus one open tail per chunk
uint64_t k12_num_tiles(uint64_t n_own) { return k12s_num_chunks(n_own); }
uint64_t k12_long_cap(uint64_t n_total) { return n_total / (k12s::MAXFAST + 1) + k12_num_tiles(n_total) + 16; }

// Records (.clrs) are emitted in two passes: a counting pass (phases A + B only) leaves the number of
// clusters of every chunk in tile_state, an exclusive scan turns it into record offsets, and the
// emitting pass (with or without scoring) reads its offset there -- no look-back between warps.
cudaError_t launch_k12(const K12Args &a, bool score, bool emit, bool exact32, cudaStream_t s)
{
    if (a.n_own == 0) return cudaSuccess;
#if LIME_K12_EXP & 4
    if ((uint64_t)a.n_reads + a.n_refs >= 0x07ffffffull) return cudaErrorInvalidValue;   // packed stream elements
#endif
    if (a.peer_on) {
        // multi-GPU peer mode: scoring only, fire-and-forget adds into the owners' slabs
        if (!score || emit || exact32 || !a.sum_t) return cudaErrorInvalidValue;
        return a.ebwt ? launch_variant<true, false, true, false, true>(a, s)
                      : launch_variant<true, false, false, false, true>(a, s);
    }
    if (emit) {
        cudaError_t e = launch_variant<false, false, false, false>(a, s);
        if (e != cudaSuccess) return e;
        k12s::chunk_scan_kernel<<<1, k12s::SCAN_THREADS, 0, s>>>(a.tile_state, k12s_num_chunks(a.n_own), nullptr);
        e = cudaGetLastError();
        if (e != cudaSuccess) return e;
        e = cudaMemsetAsync(a.n_clusters, 0, sizeof(unsigned long long), s);      // the counting pass added its total
        if (e != cudaSuccess) return e;
    }
    const bool ebwt = a.ebwt != nullptr;
#define LIME_LV(SC, EM, EB, EX) launch_variant<SC, EM, EB, EX>(a, s)
    LIME_K12S_DISPATCH(LIME_LV)
#undef LIME_LV
}

} // namespace lime
