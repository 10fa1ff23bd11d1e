// k12_stream.cu -- steps 1 and 2a in ONE pass over lcp / da / ebwt, warp-autonomous:
// alpha-cluster detection (ClusterLCP.cpp:139-283) fused with the per-cluster similarity
// accumulation (clusterAnalyze, ClusterBWT_DA.cpp:256-358).
//
// A position i with lcp[i] < alpha is a "head"; a segment is [head, next head); a segment with at
// least one da < n_reads and one da >= n_reads is an alpha-cluster.  The kernel is persistent: every
// WARP walks over chunks of 512 positions (16 words of 32) with a stride of 480, owns the segments
// whose head lies in the first 480 positions (the chunk ownership rule of ClusterLCP.cpp:150-202,
// the last word is look-ahead) and never synchronises with another warp:
//
//   stage    the next chunk of lcp / da / ebwt is copied global -> shared by the TMA engine: ONE lane
//            issues three bulk copies (cp.async.bulk, UBLKCP in SASS) against the warp's mbarrier while
//            the current chunk is scored; the warp's only DRAM reads, no load latency on the critical path.
//   phase A  lane = position inside a word: two ballots per word give the head word (lcp < alpha)
//            and the read word (da < n_reads); lane k keeps word k.
//   phase B  lane = word: bit-parallel segmented OR-fills (Kogge-Stone inside the word, the same
//            routine over the words through ballots) decide which positions lie in a segment
//            that holds both a read and a genome: the "active" word.  No per-cluster loop.
//   phase C  active positions (about a fifth of the input) are compacted from the stage into the
//            warp's shared-memory stream, one packed word id | class << 27 | cluster-start << 31 each.
//   phase D  32-lane windows slide over the stream, one lane per position: equal ids are grouped
//            with __match_any_sync, symbol classes counted with ballots, every read leader meets
//            every genome leader of its cluster by shuffle.  An update (read, genome, t) is either
//              MODE_DIRECT  a fire-and-forget add into the uint8 similarity matrix (one GPU), or
//              MODE_FLAT    an 8-byte record appended to a flat list in local HBM (multi-GPU: the records
//                           are routed to the read owner afterwards, xchg.cu).
//   Clusters longer than 32 positions or running past the look-ahead are rare; they are appended
//   to a list that the block-path kernel of k2_similarity.cu consumes afterwards.
//
// lcp is read as uint32 (the file format, Tools.h:48) or -- template LCP8 -- as the byte min(lcp, 255)
// a narrowed host load left on the device (lime_set_alpha_limit): the kernel only ever tests lcp < alpha.
//
// Algorithmic traffic: (8 + E) bytes per position, E = 1 with ebwt; + 16 B per cluster if records are emitted.
#include "common.cuh"
#include "score.cuh"
#include <algorithm>

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
constexpr uint32_t ID_MASK = K12_ID_LIMIT;   // packed stream element: id | class << 27 | start << 31
constexpr int CTAS_PER_SM = 8;            // 8 x 4 warps: 64 registers, 27.4 KB of shared memory per CTA

enum { MODE_COUNT = 0, MODE_EMIT = 1, MODE_DIRECT = 2, MODE_FLAT = 3 };

// x_i |= x_{i-1} while c_i (position i continues the run of i-1): a segmented OR-fill towards the high bits.
// One addition does it: with B = c | x the carry INTO position i of x + B is the OR of the x bits below i inside
// its run (a run start has c = 0, so it generates a carry exactly when its own x is set and kills the one that
// arrives from the run below); the carry vector of a sum is sum ^ a ^ b.  Three instructions instead of the
// fifteen of a Kogge-Stone ladder, and the add runs on the FMA pipe (the kernel is bound by the ALU pipe).
__device__ __forceinline__ uint32_t fill_up(uint32_t x, uint32_t c)
{
    const uint32_t b = c | x;
    const uint32_t carry_in = (x + b) ^ x ^ b;
    return x | (carry_in & c);
}
// x_i |= x_{i+1} while c_i (position i continues into i+1): the same on the bit-reversed words
__device__ __forceinline__ uint32_t fill_down(uint32_t x, uint32_t c)
{
    return __brev(fill_up(__brev(x), __brev(c)));
}
__device__ __forceinline__ uint32_t mask_lt(int k)   // bits [0, k), k may be <= 0 or >= 32
{
    return k <= 0 ? 0u : (k >= 32 ? 0xffffffffu : ((1u << k) - 1u));
}

template <bool LCP8> struct LcpElem { typedef uint32_t type; };
template <> struct LcpElem<true> { typedef uint8_t type; };

template <bool LCP8>
struct alignas(16) WarpSmem {
    typename LcpElem<LCP8>::type lcp[CH];   // stage of the next chunk, filled by the bulk copies while this one is
                                            // scored (dead after phase A: phase C keeps its per-word broadcast there)
    uint32_t da[CH];
    uint8_t  eb[CH];
    uint32_t st[SCAP];                      // compacted stream: id | class << 27 | start << 31
};
template <bool LCP8>
struct alignas(16) Smem {
    WarpSmem<LCP8> w[WARPS];
    unsigned long long mbar[WARPS];         // one transaction barrier per warp
    unsigned pipe[WARPS];                   // per warp: bit 0 = a staged chunk is in flight, bit 1 = its barrier phase
    uint8_t cls[256];                       // symbol -> class
};

// ---- TMA bulk copy global -> shared against an mbarrier (sm_90+) ---------------------------------------
__device__ __forceinline__ void mbar_init(unsigned mbar, unsigned count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(mbar), "r"(count) : "memory");
}
__device__ __forceinline__ void fence_mbar_init()
{
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
// generic-proxy accesses to shared memory issued so far are ordered before later async-proxy (TMA) accesses
__device__ __forceinline__ void fence_proxy_async_smem()
{
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(unsigned mbar, unsigned bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(mbar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_g2s(unsigned dst, const void *src, unsigned bytes, unsigned mbar, uint64_t pol)
{
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;"
                 ::"r"(dst), "l"(src), "r"(bytes), "r"(mbar), "l"(pol) : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned mbar, unsigned parity)
{
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "K12_WAIT_%=:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra K12_DONE_%=;\n"
        "bra K12_WAIT_%=;\n"
        "K12_DONE_%=:\n"
        "}\n" ::"r"(mbar), "r"(parity) : "memory");
}
// L2 eviction policies: the inputs are read once (evict_first) and should not push the similarity matrix,
// whose lines are re-visited by later adds (evict_last), out of L2: 2.11 -> 1.83 ms per strand in round 1
__device__ __forceinline__ uint64_t policy_evict_first()
{
    uint64_t p;
    asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(p));
    return p;
}
__device__ __forceinline__ uint64_t policy_evict_last()
{
    uint64_t p;
    asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(p));
    return p;
}
// the two policy words, once per process (they are kernel parameters afterwards: no registers held for them)
__global__ void policy_kernel(uint64_t *out)
{
    out[0] = policy_evict_first();
    out[1] = policy_evict_last();
}
__device__ __forceinline__ void red_add_hint(unsigned *addr, unsigned v, uint64_t pol)
{
    asm volatile("red.relaxed.gpu.global.add.L2::cache_hint.u32 [%0], %1, %2;" ::"l"(addr), "r"(v), "l"(pol) : "memory");
}

// four consecutive lcp values at position p (p a multiple of 4; buffers are 16-byte aligned), guarded at the end
template <bool LCP8>
__device__ __forceinline__ void load4_lcp(const void *base, uint64_t p, uint64_t n_total, uint32_t l[4])
{
    if (p + 4 <= n_total) {
        if (LCP8) {
            const uint32_t v = __ldg(reinterpret_cast<const uint32_t *>(static_cast<const uint8_t *>(base) + p));
            l[0] = v & 255u; l[1] = (v >> 8) & 255u; l[2] = (v >> 16) & 255u; l[3] = v >> 24;
        } else {
            const uint4 v = ldg128(static_cast<const uint32_t *>(base) + p);
            l[0] = v.x; l[1] = v.y; l[2] = v.z; l[3] = v.w;
        }
    } else {
#pragma unroll
        for (int e = 0; e < 4; e++) {
            const bool ok = p + e < n_total;
            if (LCP8) l[e] = ok ? (uint32_t)__ldg(static_cast<const uint8_t *>(base) + p + e) : 0xffffffffu;
            else      l[e] = ok ? __ldg(static_cast<const uint32_t *>(base) + p + e) : 0xffffffffu;
        }
    }
}

// ---- one similarity update (read, genome g, t) -----------------------------------------------------------
// MODE_DIRECT: fire-and-forget add of t << 8*(g&3) on the 32-bit word that holds the uint8 cell: no return
//   value to wait for.  A carry out of a byte (true cell sum > 255, where the reference wraps,
//   ClusterBWT_DA.cpp:183,247) lowers the byte sum of the matrix by 255 or 256, so the host proves their
//   absence afterwards by comparing the byte sum of the matrix with the total of all t (sum_t) and otherwise
//   redoes the strand exactly.
// MODE_FLAT: record {read, genome << 8 | t} appended to the warp's current block of the flat list; a block of
//   K12_FLAT_BLK slots is filled completely before the next one is taken from the global cursor.
// Called by all 32 lanes together (t = 0: nothing to add for this lane).
struct UpdState {
    unsigned updates;                 // warp-uniform: updates issued by the warp
    unsigned t_total;                 // per lane: sum of their t
    unsigned long long fl_base;       // warp-uniform: first slot of the current block
    unsigned fl_fill;                 // warp-uniform: slots used in it (K12_FLAT_BLK: none taken yet)
};
template <int MODE>
__device__ __forceinline__ void emit_update(const K12Args &P, UpdState &U, unsigned lane, uint32_t read, uint32_t g, uint32_t t,
                                            unsigned *row_w)
{
    if (MODE == MODE_DIRECT) {
        if (t) {
            red_add_hint(row_w + (g >> 2), t << ((g & 3u) * 8u), P.pol_evict_last);
            U.t_total += t;
        }
        U.updates += (unsigned)__popc(__ballot_sync(FULL, t != 0u));
    } else {
        const unsigned m = __ballot_sync(FULL, t != 0u);
        if (m) {                                                  // warp-uniform
            const unsigned n = (unsigned)__popc(m);
            const bool spill = U.fl_fill + n > (unsigned)K12_FLAT_BLK;
            unsigned long long nb = 0;
            if (spill) {
                if (lane == 0) nb = atomicAdd(P.flat_count, (unsigned long long)K12_FLAT_BLK);
                nb = __shfl_sync(FULL, nb, 0);
            }
            const unsigned idx = U.fl_fill + (unsigned)__popc(m & ((1u << lane) - 1u));
            const unsigned long long dst = idx < (unsigned)K12_FLAT_BLK ? U.fl_base + idx : nb + (idx - K12_FLAT_BLK);
            if (t) {
                if (dst < P.flat_cap) P.flat[dst] = make_uint2(read, (g << 8) | t);
                U.t_total += t;
            }
            U.updates += n;
            if (spill) { U.fl_base = nb; U.fl_fill = U.fl_fill + n - K12_FLAT_BLK; }
            else U.fl_fill += n;
        }
    }
}

// A window that holds a symbol class beyond A, C, G, T (N, IUPAC codes, the end marker: about one window in
// a hundred): exact per-class counts of all 16 classes and the full matcher.  Kept out of line so that its
// registers (two 16-byte count vectors, the matcher's arrays) do not weigh on the common path.
template <int MODE>
__device__ __noinline__ void wide_window(const K12Args &P, UpdState &U, uint32_t id, uint32_t cls, unsigned mm, unsigned cm,
                                         bool leader)
{
    const unsigned lane = lane_id();
    const bool real = id != ID_MASK, is_read = id < P.n_reads;
    unsigned present = __reduce_or_sync(FULL, real ? (1u << cls) : 0u);
    const bool iupac = (present & 0x7ff0u) != 0;
    uint4 vec = make_uint4(0, 0, 0, 0);
    while (present) {
        const int cc = __ffs(present) - 1;
        present &= present - 1;
        const unsigned b = __ballot_sync(FULL, real && cls == (uint32_t)cc);
        const unsigned v = (unsigned)__popc(mm & b) << ((cc & 3) * 8);
        if ((cc >> 2) == 0) vec.x |= v; else if ((cc >> 2) == 1) vec.y |= v;
        else if ((cc >> 2) == 2) vec.z |= v; else vec.w |= v;
    }
    unsigned gl = __ballot_sync(FULL, leader && !is_read) & cm;
    if (!(leader && is_read)) gl = 0;
    unsigned *row_w = nullptr;
    if (MODE == MODE_DIRECT) row_w = reinterpret_cast<unsigned *>(P.sim8 + (size_t)(gl ? id : 0u) * P.pitch8);
    while (__any_sync(FULL, gl != 0)) {
        const int src = gl ? 31 - __clz(gl) : 0;
        uint4 Gv;
        Gv.x = __shfl_sync(FULL, vec.x, src); Gv.y = __shfl_sync(FULL, vec.y, src);
        Gv.z = __shfl_sync(FULL, vec.z, src); Gv.w = __shfl_sync(FULL, vec.w, src);
        const uint32_t g = __shfl_sync(FULL, id, src) - P.n_reads;
        uint32_t t = (iupac ? score_iupac(vec, Gv) : score_plain(vec, Gv)) & 0xffu;
        if (gl == 0) t = 0;
        else gl ^= 1u << src;
        emit_update<MODE>(P, U, lane, id, g, t, row_w);
    }
}

template <int MODE, bool EBWT, bool LCP8>
__global__ void __launch_bounds__(THREADS, CTAS_PER_SM) stream_kernel(const __grid_constant__ K12Args P)
{
    constexpr bool SCORE = MODE >= MODE_DIRECT;
    constexpr bool EMIT = MODE == MODE_EMIT;
    typedef typename LcpElem<LCP8>::type lcp_t;
    __shared__ Smem<LCP8> sm;
    // the warp index through a shuffle: the compiler then knows that it, the chunk counters and the stage addresses
    // derived from it are warp-uniform and keeps them in uniform registers (bulk copies take uniform operands)
    const unsigned lane = lane_id(), warp = __shfl_sync(FULL, threadIdx.x >> 5, 0);
    const uint32_t lt_mask = (1u << lane) - 1u;
    if (SCORE && EBWT) {
        for (int i = threadIdx.x; i < 256; i += THREADS) sm.cls[i] = d_class_table.v[i];
    }
    WarpSmem<LCP8> &ws = sm.w[warp];
    const unsigned mb = (unsigned)__cvta_generic_to_shared(&sm.mbar[warp]);
    if (lane == 0) { mbar_init(mb, 1u); fence_mbar_init(); }
    __syncthreads();
    const int w = (int)lane;                                  // the word this lane owns in phase B

    const uint32_t n_chunks = (uint32_t)((P.n_own + OWN - 1) / OWN);       // < 2^32: checked by the launcher
    // warp-uniform running totals (the compiler keeps them in uniform registers, off the vector register file)
    unsigned cl_count = 0;                                    // clusters owned by this warp's chunks
    unsigned max_len = 0;                                     // longest cluster seen by this warp (open tails apart)
    // updates issued / sum of their t per lane (32 bits: a lane adds < 2^32 for every input the launcher
    // accepts), and in flat mode the warp's current block of the record list
    UpdState U;
    U.updates = 0; U.t_total = 0; U.fl_base = 0; U.fl_fill = K12_FLAT_BLK;
    const lcp_t *g_lcp = static_cast<const lcp_t *>(P.lcp);

    // Stage pipeline: the chunk after the current one is copied global -> shared by three bulk copies
    // while the current one is scored; whole chunks only, the tail of the buffer and unaligned buffers
    // are loaded synchronously.
    const bool aligned = ((reinterpret_cast<uintptr_t>(P.lcp) | reinterpret_cast<uintptr_t>(P.da) |
                           (SCORE && EBWT ? reinterpret_cast<uintptr_t>(P.ebwt) : 0)) & 15u) == 0;
    constexpr unsigned STAGE_BYTES = CH * sizeof(lcp_t) + CH * 4u + (SCORE && EBWT ? (unsigned)CH : 0u);
    auto stage_issue = [&](uint64_t cc) -> bool {             // all lanes, after a __syncwarp: the stage is free
        if (cc >= n_chunks) return false;
        const uint64_t aa = cc * OWN;
        if (!aligned || aa + CH > P.n_total) return false;
        if (lane == 0) {
            const uint64_t pol_in = P.pol_evict_first;
            fence_proxy_async_smem();
            mbar_expect_tx(mb, STAGE_BYTES);
            bulk_g2s((unsigned)__cvta_generic_to_shared(ws.lcp), g_lcp + aa, CH * sizeof(lcp_t), mb, pol_in);
            bulk_g2s((unsigned)__cvta_generic_to_shared(ws.da), P.da + aa, CH * 4u, mb, pol_in);
            if (SCORE && EBWT) bulk_g2s((unsigned)__cvta_generic_to_shared(ws.eb), P.ebwt + aa, CH, mb, pol_in);
        }
        return true;
    };
    // Chunk schedule: a grid stride, warp by warp.  (One contiguous range per CTA was measured too: 1.93 instead of
    // 1.84 ms; so was pulling the chunk after the next one into L2 ahead of its copy, by bulk prefetch or by
    // per-lane prefetch instructions: 2.0-2.1 ms; and per-lane cp.async instead of the bulk copies: the same
    // 1.84 ms.  What bounds the kernel is the time a chunk's copy takes to arrive while the random adds into the
    // 224 MB matrix miss L2: with the adds confined to an L2-resident 7 MB it runs at 1.41 ms.  profiles/README.md)
    auto chunk_of = [&](uint32_t t) -> uint64_t {
        return (uint64_t)blockIdx.x * WARPS + warp + (uint64_t)t * gridDim.x * WARPS;
    };
    const uint64_t c_first = chunk_of(0);
    // whether a staged chunk is in flight and the phase of its barrier live in shared memory, not in registers
    unsigned &pipe = sm.pipe[warp];
    {
        const bool st0 = stage_issue(c_first);
        if (lane == 0) pipe = st0 ? 1u : 0u;
        __syncwarp();
    }

    uint32_t t_it = 0;
    for (uint64_t c = c_first; c < n_chunks; c = chunk_of(++t_it)) {
        const uint64_t a = c * OWN;                                  // buffer-local start of the chunk
        const int Lt = (int)min((uint64_t)OWN, P.n_own - a);         // owned positions
        const int Lext = (int)min((uint64_t)CH, P.n_total - a);      // real positions in the chunk

        // ---- phase A: head / read words (lane = position inside a word) ------------------------------
        uint32_t H = 0xffffffffu, R = 0;                             // lanes >= WORDS: beyond the chunk
        const unsigned pst = pipe;
        if (pst & 1u) {
            mbar_wait(mb, pst >> 1);
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
            const lcp_t *pl = g_lcp + a + lane;
            const uint32_t *pd = P.da + a + lane;
            const uint8_t *pe = P.ebwt + a + lane;
#pragma unroll
            for (int k = 0; k < WORDS; k++) {
                const bool ok = 32 * k + (int)lane < Lext;
                const uint32_t l = ok ? (uint32_t)__ldcs(pl + 32 * k) : 0u;    // virtual head
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
        auto issue_next = [&]() {
            __syncwarp();
            const bool st = stage_issue(chunk_of(t_it + 1));
            // the phase bit flips whenever a staged chunk has been consumed
            if (lane == 0) pipe = (st ? 1u : 0u) | ((((pst >> 1) ^ pst) & 1u) << 1);
            __syncwarp();
        };
        if (!SCORE) issue_next();

        // ---- phase B: per-word segment flags (lane w owns word w) -------------------------------------
        // A segment holds a read and a genome iff two neighbours inside it differ in kind:
        // X_i = (read_i != read_{i+1}) and i+1 is not a head.  One flag to spread over the segment.
        const uint32_t V = mask_lt(Lext - w * 32);
        const uint32_t Hup = __shfl_down_sync(FULL, H, 1), Rup = __shfl_down_sync(FULL, R, 1);
        const uint32_t Hn = (H >> 1) | (Hup << 31), Rn = (R >> 1) | (Rup << 31);
        const uint32_t X = (R ^ Rn) & ~Hn & mask_lt(Lext - w * 32 - 1);   // the last position has no neighbour
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
            const uint32_t below_lt = mask_lt(Lt - w * 32);
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
                load4_lcp<LCP8>(P.lcp, p, P.n_total, l);
                if (p + 4 <= P.n_total) {
                    const uint4 dv = ldg128(P.da + p);
                    d[0] = dv.x; d[1] = dv.y; d[2] = dv.z; d[3] = dv.w;
                } else {
#pragma unroll
                    for (int e = 0; e < 4; e++) d[e] = p + e < P.n_total ? __ldg(P.da + p + e) : 0u;
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
        const unsigned chunk_count = __reduce_add_sync(FULL, count);
        cl_count += chunk_count;
        if (own_tail) atomicMax(P.max_len, tail_len);                     // rare; may exceed 32 bits

        if (MODE == MODE_COUNT) {
            // counting pass of the two-pass record emission: clusters owned by this chunk
            if (lane == 0) P.tile_state[c] = chunk_count;
            continue;
        }
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
            unsigned long long rec_base = e0 + (incl - count);
            // every owned head: length to the next head, record in ascending pStart order
            uint32_t hb = heads;
            unsigned lane_max = 0;
            while (hb) {
                const int b = __ffs(hb) - 1;
                hb &= hb - 1;
                const uint32_t m = H & ~((2u << b) - 1u);              // next head in this word (b < 31 or m == 0)
                const int j = (b < 31 && m) ? w * 32 + __ffs(m) - 1 : nh_above;
                const int len = j - (w * 32 + b);
                lane_max = max(lane_max, (unsigned)len);
                if (rec_base < P.out_cap)
                    *reinterpret_cast<ulonglong2 *>(&P.out[rec_base]) = make_ulonglong2(P.pos_base + a + w * 32 + b, (unsigned long long)len);
                rec_base++;
            }
            if (own_tail && rec_base < P.out_cap)
                *reinterpret_cast<ulonglong2 *>(&P.out[rec_base]) = make_ulonglong2(P.pos_base + a + hl, tail_len);
            max_len = max(max_len, __reduce_max_sync(FULL, lane_max));
            continue;
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
                max_len = max(max_len, (unsigned)(j - i));
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

        // ---- phase C: compact the active positions into the warp's stream ------------------------------
        // lane = position inside a word again: per word one broadcast of (active, head, offset), the
        // active lanes copy id / symbol class / start flag from the stage; words without active positions
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
#pragma unroll
                for (int k = 0; k < WORDS; k++) {
                    const uint4 inf = info[k];
                    if (inf.x == 0u) continue;
                    if ((inf.x >> lane) & 1u) {
                        const uint32_t o = inf.z + __popc(inf.x & lt_mask);
                        const uint32_t cl0 = EBWT ? (uint32_t)sm.cls[ws.eb[32 * k + lane]] : 0u;
                        ws.st[o] = ws.da[32 * k + lane] | (cl0 << 27) | (((inf.y >> lane) & 1u) << 31);
                    }
                }
            }
        }
        issue_next();                                                     // includes the __syncwarp
        if (n_act == 0) continue;                                         // warp-uniform
        ws.st[n_act + lane] = 0xffffffffu;                                // one window of padding: lone starts
        __syncwarp();

        // ---- phase D: score the clusters of the stream, 32-lane windows --------------------------------
        unsigned longest = 0;                                             // warp-uniform
        // software pipeline over the windows: the elements of the next window are loaded while this one is scored
        int i = 0;
        uint32_t pk = ws.st[lane], nx = ws.st[32];
        while (i < n_act) {
            // equal ids of the window; restricted to the lanes of the own cluster further down.  (Issued first: the
            // match is the longest-latency operation of the window set-up.)
            const unsigned mraw = __match_any_sync(FULL, pk & ID_MASK);
            uint32_t id = pk & ID_MASK;                                   // ID_MASK itself: padding
            uint32_t sb = __ballot_sync(FULL, (pk >> 31) != 0u);          // bit 0 is set: i is a start
            if (!(nx >> 31)) {
                // the last cluster of the window is cut: leave it to the next window (turn its lanes
                // into lone starts; a cluster has at most 32 positions, so it does not start at lane 0)
                const int ls = 31 - __clz(sb);
                sb |= 0xffffffffu << ls;
                if ((int)lane >= ls) id = ID_MASK;
                i += ls;
            } else {
                i += 32;
            }
            const uint32_t cur = pk;
            {
                const bool more = i < n_act;                              // warp-uniform
                pk = more ? ws.st[i + lane] : 0xffffffffu;
                nx = more ? ws.st[i + 32] : 0xffffffffu;
            }
            const uint32_t le_mask = lt_mask | (1u << lane);
            const uint32_t start_bit = 0x80000000u >> __clz(sb & le_mask);     // first lane of my cluster
            const uint32_t later = sb & ~le_mask;
            const uint32_t upper = (later & (0u - later)) - 1u;                // lanes below the next start (all if none)
            const uint32_t cm = upper & ~(start_bit - 1u);                     // lanes of my cluster
            const bool is_read = id < P.n_reads;
            const unsigned mm = mraw & cm;                                     // lanes with my id in my cluster
            const bool leader = (mm & lt_mask) == 0;
            longest = max(longest, __reduce_max_sync(FULL, (unsigned)__popc(cm)));

            // per-class counts of my id: popc(my group & lanes of that class), A / C / G / T in the four bytes
            uint32_t vec = 0;
            if (!EBWT) {
                vec = (unsigned)__popc(mm);
            } else {
                const uint32_t cls = (cur >> 27) & 15u;
                const unsigned rest = __ballot_sync(FULL, cls >= 4u && id != ID_MASK);
                if (rest) {                                                    // a class beyond A,C,G,T (warp-uniform, rare)
                    UpdState W = U;
                    wide_window<MODE>(P, W, id, cls, mm, cm, leader);
                    U = W;
                    continue;
                }
                const unsigned c0 = __ballot_sync(FULL, (cls & 1u) != 0), c1 = __ballot_sync(FULL, (cls & 2u) != 0);
                vec = (unsigned)__popc(mm & ~c1 & ~c0) | ((unsigned)__popc(mm & ~c1 & c0) << 8) |
                      ((unsigned)__popc(mm & c1 & ~c0) << 16) | ((unsigned)__popc(mm & c1 & c0) << 24);
            }
            // reads wrap at 256 (:123,206), genomes saturate at 255 (:96-97,222-223): counts <= 32 here

            // every read leader against every genome leader of its cluster
            unsigned gl = __ballot_sync(FULL, leader && !is_read) & cm;
            if (!(leader && is_read)) gl = 0;
            // the row of a read leader, once per window
            unsigned *row_w = nullptr;
            if (MODE == MODE_DIRECT) row_w = reinterpret_cast<unsigned *>(P.sim8 + (size_t)(gl ? id : 0u) * P.pitch8);
            while (__any_sync(FULL, gl != 0)) {
                const int src = gl ? 31 - __clz(gl) : 0;
                const uint32_t Gx = __shfl_sync(FULL, vec, src);
                const uint32_t g = __shfl_sync(FULL, id, src) - P.n_reads;
                uint32_t t = EBWT ? score_acgt(vec, Gx) : min(vec, Gx);
                if (gl == 0) t = 0;
                else gl ^= 1u << src;
                emit_update<MODE>(P, U, lane, id, g, t, row_w);
            }
        }
        max_len = max(max_len, longest);
    }
    if (MODE == MODE_FLAT && U.fl_fill < (unsigned)K12_FLAT_BLK) {
        // null records (t = 0) up to the end of the warp's last block: every slot below flat_count is defined
        for (unsigned i = U.fl_fill + lane; i < (unsigned)K12_FLAT_BLK; i += 32)
            if (U.fl_base + i < P.flat_cap) P.flat[U.fl_base + i] = make_uint2(0xffffffffu, 0u);
    }
    // ---- totals of this warp: one atomic per warp and scalar -------------------------------------------
    {
        const unsigned c2 = cl_count, mx = max_len;
        const unsigned long long up = U.updates;
        unsigned long long tt = U.t_total;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) tt += __shfl_xor_sync(FULL, tt, o);
        if (lane == 0) {
            if (!EMIT && c2) atomicAdd(P.n_clusters, (unsigned long long)c2);
            if (mx) atomicMax(P.max_len, (unsigned long long)mx);
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

} // namespace k12s

uint64_t k12s_num_chunks(uint64_t n_own) { return (n_own + k12s::OWN - 1) / k12s::OWN; }

template <int MODE, bool EB, bool L8>
static unsigned grid_of(uint64_t n_own, int max_ctas_per_sm)
{
    static int ctas_per_sm = 0, sms = 0;
    if (!ctas_per_sm) {
        int dev = 0;
        if (cudaGetDevice(&dev) != cudaSuccess) return 0;
        if (cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess) return 0;
        if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&ctas_per_sm, k12s::stream_kernel<MODE, EB, L8>, k12s::THREADS, 0) != cudaSuccess) return 0;
        if (ctas_per_sm < 1) ctas_per_sm = 1;
    }
    const uint64_t chunks = k12s_num_chunks(n_own);
    const uint64_t want = (chunks + k12s::WARPS - 1) / k12s::WARPS;          // at least one chunk per warp
    const int per_sm = max_ctas_per_sm > 0 ? std::min(max_ctas_per_sm, ctas_per_sm) : ctas_per_sm;
    return (unsigned)std::min<uint64_t>(want, (uint64_t)sms * per_sm);
}

template <int MODE, bool EB, bool L8>
static cudaError_t launch_variant(const K12Args &a, cudaStream_t s)
{
    // persistent grid, CTAs stride over groups of chunks
    const unsigned grid = grid_of<MODE, EB, L8>(a.n_own, a.max_ctas_per_sm);
    if (!grid) return cudaErrorUnknown;
    k12s::stream_kernel<MODE, EB, L8><<<grid, k12s::THREADS, 0, s>>>(a);
    return cudaGetLastError();
}

template <int MODE, bool EB>
static cudaError_t launch_lcp(const K12Args &a, cudaStream_t s)
{
    return a.lcp8 ? launch_variant<MODE, EB, true>(a, s) : launch_variant<MODE, EB, false>(a, s);
}

// per-chunk cluster counts / record offsets are per chunk; the long-cluster list holds at most one record per
// 33 positions plus one open tail per chunk
uint64_t k12_num_tiles(uint64_t n_own) { return k12s_num_chunks(n_own); }
uint64_t k12_long_cap(uint64_t n_total) { return n_total / (k12s::MAXFAST + 1) + k12_num_tiles(n_total) + 16; }

// what: K12_DETECT  records (.clrs) in two passes: a counting pass (phases A + B only) leaves the number of
//                   clusters of every chunk in tile_state, an exclusive scan turns it into record offsets,
//                   the emitting pass reads its offset there -- no look-back between warps;
//       K12_COUNT   nClusters / maxLen only (the counting pass alone);
//       K12_DIRECT / K12_FLAT  score, see the head of this file.
cudaError_t k12_policies(uint64_t out[2])
{
    uint64_t *d = nullptr;
    cudaError_t e = cudaMalloc(&d, 16);
    if (e != cudaSuccess) return e;
    k12s::policy_kernel<<<1, 1>>>(d);
    e = cudaMemcpy(out, d, 16, cudaMemcpyDeviceToHost);
    cudaFree(d);
    return e;
}

cudaError_t launch_k12(const K12Args &a, int what, cudaStream_t s)
{
    if (a.n_own == 0) return cudaSuccess;
    if (k12s_num_chunks(a.n_own) >= 0xffffffffull) return cudaErrorInvalidValue;
    if (a.lcp8 && a.alpha > 255) return cudaErrorInvalidValue;
    if (what == K12_COUNT) return launch_lcp<k12s::MODE_COUNT, false>(a, s);
    if (what == K12_DETECT) {
        cudaError_t e = launch_lcp<k12s::MODE_COUNT, false>(a, s);
        if (e != cudaSuccess) return e;
        k12s::chunk_scan_kernel<<<1, k12s::SCAN_THREADS, 0, s>>>(a.tile_state, k12s_num_chunks(a.n_own), nullptr);
        e = cudaGetLastError();
        if (e != cudaSuccess) return e;
        // the counting pass added its totals: the emitting pass writes nClusters itself and finds the same maxLen
        e = cudaMemsetAsync(a.n_clusters, 0, sizeof(unsigned long long), s);
        if (e != cudaSuccess) return e;
        return launch_lcp<k12s::MODE_EMIT, false>(a, s);
    }
    // packed stream elements: ids below K12_ID_LIMIT (the ABI routes larger collections through the cluster list)
    if ((uint64_t)a.n_reads + a.n_refs >= K12_ID_LIMIT) return cudaErrorInvalidValue;
    const bool ebwt = a.ebwt != nullptr;
    if (what == K12_DIRECT) return ebwt ? launch_lcp<k12s::MODE_DIRECT, true>(a, s) : launch_lcp<k12s::MODE_DIRECT, false>(a, s);
    if (what == K12_FLAT) {
        if (a.n_refs >= (1u << 24)) return cudaErrorInvalidValue;        // record = genome << 8 | t
        return ebwt ? launch_lcp<k12s::MODE_FLAT, true>(a, s) : launch_lcp<k12s::MODE_FLAT, false>(a, s);
    }
    return cudaErrorInvalidValue;
}

} // namespace lime
