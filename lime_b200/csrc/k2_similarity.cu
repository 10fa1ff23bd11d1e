// k2_similarity.cu -- step 2a, per-cluster read x genome similarity accumulation
// (replaces clusterAnalyze, ClusterBWT_DA.cpp:256-358, with Update_ref_symb :81-105 and
// Analysis_and_updating :107-190 (EBWT=1) / :192-252 (EBWT=0)).
//
// For every cluster, positions are grouped by document id; a read r gets per-class symbol counts
// R[16] (uint8, wrapping, :123), a genome g gets G[16] (uint8, saturating at 255, :96-97); every
// (r, g) present in the cluster adds t(R, G) to SimArray_[r][g] (uint8, wrapping, :183).  The
// reference sorts each cluster to group equal ids (:316-322); only the grouping matters, so here
//   * clusters of <= 32 positions (the bulk) are packed several per warp, one lane per position,
//     and grouped with __match_any_sync on (cluster, id); class counts come from ballots + popc;
//   * longer clusters (up to sizeMaxBuf = 65536) take one CTA each and are grouped with a hash
//     table in an L2-resident scratch area.
// EBWT=0 is the same computation with every symbol in class 0 (t = min(#r, #g)).
//
// The accumulator is a dense [n_reads][pitch] matrix of uint8 cells updated with 32-bit atomics
// on the containing word.  An add that carries out of its byte would corrupt the neighbour, so
// every add inspects the returned old word and raises a flag; the host then reruns the pass into
// exact 32-bit cells (reduced mod 256 afterwards), which is the reference's wrap semantics.
#include "common.cuh"
#include "score.cuh"

namespace lime {
namespace k2 {

template <bool EXACT32>
__device__ __forceinline__ void sim_add(const K2Args &P, uint32_t r, uint32_t g, uint32_t t)
{
    SimAcc acc;
    acc.sim8 = P.sim8; acc.pitch8 = P.pitch8; acc.sim32 = P.sim32; acc.lo = P.lo; acc.hi = P.hi;
    acc.n_refs = P.n_refs; acc.carry = P.carry;
    lime::sim_add<EXACT32>(acc, r, g, t);
}

// ---- warp path: clusters of <= 32 positions ---------------------------------------------------
constexpr int WARP_THREADS = 256;
constexpr int WARPS = WARP_THREADS / 32;

template <bool EBWT, bool EXACT32>
__global__ void __launch_bounds__(WARP_THREADS) warp_kernel(K2Args P)
{
    __shared__ uint8_t s_owner[WARPS][32];
    const unsigned lane = lane_id(), warp = threadIdx.x >> 5;
    const uint64_t n_chunks = (P.n_cl + 31) / 32;
    const uint64_t warp_global = (uint64_t)blockIdx.x * WARPS + warp;
    const uint64_t warp_stride = (uint64_t)gridDim.x * WARPS;
    unsigned long long updates = 0;

    for (uint64_t chunk = warp_global; chunk < n_chunks; chunk += warp_stride) {
        // one cluster descriptor per lane
        const uint64_t ci = chunk * 32 + lane;
        uint64_t lstart = 0;
        uint32_t len = 0;
        if (ci < P.n_cl) {
            const ulonglong2 c = *reinterpret_cast<const ulonglong2 *>(&P.cl[ci]);
            lstart = c.x - P.pos_base;
            len = c.y > 0xffffffffull ? 0xffffffffu : (uint32_t)c.y;
        }
        const bool large = len > 32;
        {   // hand long clusters to the block path
            const unsigned lm = __ballot_sync(FULL, large);
            if (lm) {
                unsigned long long b = 0;
                if (lane == 0) b = atomicAdd(P.large_count, (unsigned long long)__popc(lm));
                b = __shfl_sync(FULL, b, 0);
                if (large) P.large_list[b + __popc(lm & ((1u << lane) - 1u))] = ci;
            }
        }
        const uint32_t eff = large ? 0u : len;
        uint32_t incl = eff;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t n = __shfl_up_sync(FULL, incl, o);
            if (lane >= (unsigned)o) incl += n;
        }
        const uint32_t excl = incl - eff;

        // rounds: greedily take the next clusters whose positions fit in 32 lanes
        int k0 = 0;
        uint32_t done = 0;
        while (k0 < 32) {
            const unsigned fit = __ballot_sync(FULL, (int)lane >= k0 && incl - done <= 32u);
            const int k1 = k0 + __popc(fit);               // fit is a contiguous run starting at k0
            const uint32_t total = __shfl_sync(FULL, incl, k1 - 1) - done;
            if (total > 0) {
                const bool in_round = (int)lane >= k0 && (int)lane < k1 && eff > 0;
                const uint32_t rel = excl - done;           // first lane slot of my cluster
                const unsigned startmask = __reduce_or_sync(FULL, in_round ? (1u << rel) : 0u);
                if (in_round) s_owner[warp][rel] = (uint8_t)lane;
                __syncwarp();
                const bool act = lane < total;
                const unsigned below = startmask & (0xffffffffu >> (31 - lane));
                const int st = act ? 31 - __clz(below) : 0;
                const int own = act ? s_owner[warp][st] : 0;
                __syncwarp();
                const uint64_t cstart = __shfl_sync(FULL, lstart, own);
                const uint32_t clen = __shfl_sync(FULL, eff, own);

                uint32_t id = 0, cls = 0;
                if (act) {
                    const uint64_t p = cstart + (lane - st);
                    id = __ldg(P.da + p);
                    if (EBWT) cls = class_of(__ldg(P.ebwt + p));
                }
                const bool is_read = id < P.n_reads;
                const unsigned long long key =
                    act ? (((unsigned long long)own << 32) | id) : ((0xffull << 40) | lane);
                const unsigned mm = __match_any_sync(FULL, key);   // lanes with my (cluster, id)
                const bool leader = act && (__ffs(mm) - 1 == (int)lane);

                // per-class counts of my id: popc(my group & lanes of that class)
                uint4 vec = make_uint4(0, 0, 0, 0);
                unsigned present = __reduce_or_sync(FULL, act ? (1u << cls) : 0u);
                const bool iupac = (present & 0x7ff0u) != 0;
                while (present) {
                    const int c = __ffs(present) - 1;
                    present &= present - 1;
                    const unsigned b = __ballot_sync(FULL, act && cls == (uint32_t)c);
                    const unsigned v = (unsigned)__popc(mm & b) << ((c & 3) * 8);
                    if ((c >> 2) == 0) vec.x |= v; else if ((c >> 2) == 1) vec.y |= v;
                    else if ((c >> 2) == 2) vec.z |= v; else vec.w |= v;
                }

                // every read leader against every genome leader of its cluster
                const unsigned cm = (clen >= 32 ? 0xffffffffu : ((1u << clen) - 1u)) << st;
                unsigned gl = __ballot_sync(FULL, leader && !is_read) & cm;
                if (!(leader && is_read)) gl = 0;
                while (__any_sync(FULL, gl != 0)) {
                    const int src = gl ? __ffs(gl) - 1 : 0;
                    uint4 G;
                    G.x = __shfl_sync(FULL, vec.x, src); G.y = __shfl_sync(FULL, vec.y, src);
                    G.z = __shfl_sync(FULL, vec.z, src); G.w = __shfl_sync(FULL, vec.w, src);
                    const uint32_t gid = __shfl_sync(FULL, id, src);
                    if (gl) {
                        const uint32_t t = (iupac ? score_iupac(vec, G) : score_plain(vec, G)) & 0xffu;
                        if (t) { sim_add<EXACT32>(P, id, gid - P.n_reads, t); updates++; }
                        gl &= gl - 1;
                    }
                }
            }
            done += total;
            k0 = k1;
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) updates += __shfl_xor_sync(FULL, updates, o);
    if (lane == 0 && updates) atomicAdd(P.n_updates, updates);
}

// ---- block path: clusters of 33 .. 65536 positions ---------------------------------------------
// One CTA per cluster.  Positions are grouped by document id in a hash table in the CTA's (L2-resident) scratch
// area; the table is then folded into two compact arrays -- reads {id, 16 class counts}, genomes likewise -- and
// the pair phase runs out of shared memory: the genomes of the cluster (a tile of GT at a time) are staged there
// once, a warp takes a read (its record is fetched one iteration ahead) and its lanes the genomes.  Nothing in
// the pair loop waits on global memory: at alpha = 12 a strand holds hundreds of clusters of ~40,000 positions
// (~18,000 reads x all genomes each), which an earlier version with three dependent L2 loads per pair, an atomic
// whose return value was inspected and one global counter update per pair took 660 ms per strand to score.
// RED: updates are fire-and-forget adds and the caller proves the absence of byte carries by the byte sum of the
// matrix (the fused path, K12Args::sum_t); otherwise every add inspects the word it returns.
constexpr int BLOCK_THREADS = 256;
constexpr int BLOCK_WARPS = BLOCK_THREADS / 32;
constexpr int GT = 1024;                             // genomes per shared-memory tile
constexpr uint32_t HCAP = 2 * LIME_SIZE_MAX_BUF;     // slots per CTA table
constexpr uint32_t EMPTY = 0xffffffffu;
// per-CTA scratch layout: keys[HCAP] u32 | cnt[HCAP][8] u32 (16 x u16 per slot) | rid[65536] u32 | gid[65536] u32
//                         | rcnt[65536] uint4 (16 x u8) | gcnt[65536] uint4
constexpr uint64_t SCRATCH_BYTES =
    (uint64_t)HCAP * 4 + (uint64_t)HCAP * 32 + 2ull * LIME_SIZE_MAX_BUF * 4 + 2ull * LIME_SIZE_MAX_BUF * 16;

__device__ __forceinline__ uint32_t hash32(uint32_t x)
{
    x ^= x >> 16; x *= 0x7feb352du; x ^= x >> 15; x *= 0x846ca68bu; x ^= x >> 16;
    return x;
}

// a warp's cursor into the record list (multi-GPU): blocks of K12_FLAT_BLK slots, as the fused kernel takes them
struct FlatCursor { unsigned long long base; unsigned fill; };
__device__ __forceinline__ void flat_append_block(const FlatOut &F, FlatCursor &C, unsigned lane, uint32_t r, uint32_t g, uint32_t t)
{
    const unsigned m = __ballot_sync(FULL, t != 0u);
    if (m == 0u) return;
    const unsigned n = (unsigned)__popc(m);
    const bool spill = C.fill + n > (unsigned)K12_FLAT_BLK;
    unsigned long long nb = 0;
    if (spill) {
        if (lane == 0) nb = atomicAdd(F.count, (unsigned long long)K12_FLAT_BLK);
        nb = __shfl_sync(FULL, nb, 0);
    }
    const unsigned idx = C.fill + (unsigned)__popc(m & ((1u << lane) - 1u));
    const unsigned long long dst = idx < (unsigned)K12_FLAT_BLK ? C.base + idx : nb + (idx - K12_FLAT_BLK);
    if (t != 0u && dst < F.cap) F.rec[dst] = make_uint2(r, (g << 8) | t);
    if (spill) { C.base = nb; C.fill = C.fill + n - K12_FLAT_BLK; }
    else C.fill += n;
}

template <bool EBWT, bool EXACT32, bool RED>
__global__ void __launch_bounds__(BLOCK_THREADS) block_kernel(K2Args P)
{
    __shared__ unsigned long long s_item;
    __shared__ unsigned s_nr, s_ng, s_iupac;
    __shared__ uint4 s_gcnt[GT];
    __shared__ uint32_t s_gid[GT];
    uint8_t *my = P.scratch + (size_t)blockIdx.x * P.scratch_per_cta;
    uint32_t *keys = reinterpret_cast<uint32_t *>(my);
    uint32_t *cnt = keys + HCAP;                       // [HCAP][8]
    uint32_t *rid = cnt + (size_t)HCAP * 8;
    uint32_t *gid = rid + LIME_SIZE_MAX_BUF;
    uint4 *rcnt = reinterpret_cast<uint4 *>(gid + LIME_SIZE_MAX_BUF);
    uint4 *gcnt = rcnt + LIME_SIZE_MAX_BUF;
    const int tid = threadIdx.x;
    const unsigned lane = lane_id(), warp = tid >> 5;
    // the list may have overflowed (count > cap) and a cluster may exceed the buffers: the host
    // reports both as errors after the pass, the kernel just stays inside its arrays
    const unsigned long long n_large = min(*P.large_count, (unsigned long long)P.large_cap);
    unsigned long long updates = 0, tsum = 0;
    FlatCursor fc; fc.base = 0; fc.fill = K12_FLAT_BLK;
    const bool flat = P.flat.rec != nullptr;

    while (true) {
        if (tid == 0) { s_item = atomicAdd(P.large_next, 1ull); s_nr = 0; s_ng = 0; s_iupac = 0; }
        __syncthreads();
        const unsigned long long item = s_item;
        if (item >= n_large) break;
        const lime_cluster c = P.long_recs ? P.long_recs[item] : P.cl[P.large_list[item]];
        const uint64_t start = c.pStart - P.pos_base;
        if (c.len > LIME_SIZE_MAX_BUF) { __syncthreads(); continue; }    // LIME_ERR_MAXLEN on the host
        const uint32_t len = (uint32_t)c.len;
        uint32_t cap = 64;
        while (cap < 2 * len) cap <<= 1;
        const uint32_t mask = cap - 1;

        for (uint32_t h = tid; h < cap; h += BLOCK_THREADS) keys[h] = EMPTY;
        {
            uint4 *z = reinterpret_cast<uint4 *>(cnt);
            for (uint32_t h = tid; h < cap * 2; h += BLOCK_THREADS) z[h] = make_uint4(0, 0, 0, 0);
        }
        __syncthreads();

        // group by id: per-(id, class) counts as 16 x u16
        for (uint32_t e = tid; e < len; e += BLOCK_THREADS) {
            const uint32_t id = __ldg(P.da + start + e);
            const uint32_t cls = EBWT ? class_of(__ldg(P.ebwt + start + e)) : 0u;
            uint32_t h = hash32(id) & mask;
            while (true) {
                uint32_t k = __ldcg(&keys[h]);
                if (k == EMPTY) {
                    k = atomicCAS(&keys[h], EMPTY, id);
                    if (k == EMPTY) break;          // slot claimed for id
                }
                if (k == id) break;
                h = (h + 1) & mask;
            }
            atomicAdd(&cnt[(size_t)h * 8 + (cls >> 1)], 1u << ((cls & 1) * 16));
            if (cls >= 4 && cls <= 14) s_iupac = 1;
        }
        __syncthreads();

        // fold the table into the compact arrays; counts to uint8: reads wrap (:123), genomes saturate (:96-97)
        for (uint32_t h = tid; h < cap; h += BLOCK_THREADS) {
            const uint32_t id = __ldcg(&keys[h]);   // L2: the slot was written by an atomic
            if (id == EMPTY) continue;
            const bool is_read = id < P.n_reads;
            const uint4 lo = __ldcg(reinterpret_cast<const uint4 *>(&cnt[(size_t)h * 8]));
            const uint4 hi = __ldcg(reinterpret_cast<const uint4 *>(&cnt[(size_t)h * 8 + 4]));
            const uint32_t in[8] = {lo.x, lo.y, lo.z, lo.w, hi.x, hi.y, hi.z, hi.w};
            uint32_t out[4];
#pragma unroll
            for (int w = 0; w < 4; w++) {
                uint32_t packed = 0;
#pragma unroll
                for (int b = 0; b < 4; b++) {
                    const int cl = w * 4 + b;
                    uint32_t v = (in[cl >> 1] >> ((cl & 1) * 16)) & 0xffffu;
                    v = is_read ? (v & 0xffu) : min(v, 255u);
                    packed |= v << (b * 8);
                }
                out[w] = packed;
            }
            const uint4 pk = make_uint4(out[0], out[1], out[2], out[3]);
            if (is_read) { const unsigned k = atomicAdd(&s_nr, 1u); rid[k] = id; rcnt[k] = pk; }
            else         { const unsigned k = atomicAdd(&s_ng, 1u); gid[k] = id - P.n_reads; gcnt[k] = pk; }
        }
        __syncthreads();

        // reads x genomes: the genomes from shared memory, a warp per read, lanes over genomes
        const uint32_t nr = s_nr, ng = s_ng;
        const bool iupac = s_iupac != 0;
        for (uint32_t gb = 0; gb < ng; gb += GT) {
            const uint32_t gt = min((uint32_t)GT, ng - gb);
            for (uint32_t i = tid; i < gt; i += BLOCK_THREADS) { s_gcnt[i] = __ldcg(gcnt + gb + i); s_gid[i] = __ldcg(gid + gb + i); }
            __syncthreads();
            uint32_t ri = warp;
            uint4 Rn = make_uint4(0, 0, 0, 0);
            uint32_t rn = 0;
            if (ri < nr) { Rn = __ldcg(rcnt + ri); rn = __ldcg(rid + ri); }
            while (ri < nr) {
                const uint4 R = Rn;
                const uint32_t r = rn;
                const uint32_t nx = ri + BLOCK_WARPS;
                if (nx < nr) { Rn = __ldcg(rcnt + nx); rn = __ldcg(rid + nx); }      // lands while this read is scored
                unsigned *row_w = (!EXACT32 && !flat) ? reinterpret_cast<unsigned *>(P.sim8 + (size_t)r * P.pitch8) : nullptr;
                for (uint32_t g0 = 0; g0 < gt; g0 += 32) {                 // warp-uniform trip count
                    const uint32_t gi = g0 + lane;
                    uint32_t t = 0, g = 0;
                    if (gi < gt) {
                        const uint4 G = s_gcnt[gi];
                        t = (iupac ? score_iupac(R, G) : score_plain(R, G)) & 0xffu;
                        g = s_gid[gi];
                    }
                    if (flat) flat_append_block(P.flat, fc, lane, r, g, t);
                    else if (t) {
                        if (RED) {
                            asm volatile("red.relaxed.gpu.global.add.u32 [%0], %1;" ::"l"(row_w + (g >> 2)), "r"(t << ((g & 3u) * 8u)) : "memory");
                        } else sim_add<EXACT32>(P, r, g, t);
                    }
                    if (t) { updates++; tsum += t; }
                }
                ri = nx;
            }
            __syncthreads();
        }
    }
    if (flat && fc.fill < (unsigned)K12_FLAT_BLK) {
        // null records (t = 0) up to the end of the warp's last block: every slot below the cursor is defined
        for (unsigned i = fc.fill + lane; i < (unsigned)K12_FLAT_BLK; i += 32)
            if (fc.base + i < P.flat.cap) P.flat.rec[fc.base + i] = make_uint2(0xffffffffu, 0u);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) { updates += __shfl_xor_sync(FULL, updates, o); tsum += __shfl_xor_sync(FULL, tsum, o); }
    if (lane == 0 && updates) atomicAdd(P.n_updates, updates);
    if (lane == 0 && tsum && P.sum_t) atomicAdd(P.sum_t, tsum);
}

// ---- exact pass epilogue: 32-bit cells -> uint8 cells (mod 256) ---------------------------------
__global__ void narrow_kernel(const uint32_t *sim32, uint32_t lo, uint32_t hi, uint32_t n_refs,
                              uint8_t *sim8, uint64_t pitch8)
{
    const uint64_t total = (uint64_t)(hi - lo) * n_refs;
    for (uint64_t x = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; x < total;
         x += (uint64_t)gridDim.x * blockDim.x) {
        const uint32_t r = (uint32_t)(x / n_refs), g = (uint32_t)(x - (uint64_t)r * n_refs);
        sim8[(size_t)(lo + r) * pitch8 + g] = (uint8_t)(sim32[x] & 0xffu);
    }
}

} // namespace k2

uint64_t k2_block_scratch_bytes() { return k2::SCRATCH_BYTES; }
int k2_block_grid() { return 148 * 3; }       // 24 warps per SM; 7 MB of scratch per CTA

cudaError_t launch_k2_warp(const K2Args &a, bool exact32, cudaStream_t s)
{
    if (a.n_cl == 0) return cudaSuccess;
    const uint64_t chunks = (a.n_cl + 31) / 32;
    const uint64_t want = (chunks + k2::WARPS - 1) / k2::WARPS;
    const unsigned grid = (unsigned)(want < 148ull * 8 * 4 ? want : 148ull * 8 * 4);
    const bool ebwt = a.ebwt != nullptr;
    if (ebwt && !exact32)       k2::warp_kernel<true, false><<<grid, k2::WARP_THREADS, 0, s>>>(a);
    else if (ebwt && exact32)   k2::warp_kernel<true, true><<<grid, k2::WARP_THREADS, 0, s>>>(a);
    else if (!ebwt && !exact32) k2::warp_kernel<false, false><<<grid, k2::WARP_THREADS, 0, s>>>(a);
    else                        k2::warp_kernel<false, true><<<grid, k2::WARP_THREADS, 0, s>>>(a);
    return cudaGetLastError();
}

cudaError_t launch_k2_block(const K2Args &a, bool exact32, cudaStream_t s, int grid)
{
    const bool ebwt = a.ebwt != nullptr;
    // fire-and-forget adds when the caller proves the absence of byte carries by the byte sum (sum_t given)
    const bool red = !exact32 && a.sum_t != nullptr && a.flat.rec == nullptr;
    if (red) {
        if (ebwt) k2::block_kernel<true, false, true><<<grid, k2::BLOCK_THREADS, 0, s>>>(a);
        else      k2::block_kernel<false, false, true><<<grid, k2::BLOCK_THREADS, 0, s>>>(a);
    }
    else if (ebwt && !exact32)  k2::block_kernel<true, false, false><<<grid, k2::BLOCK_THREADS, 0, s>>>(a);
    else if (ebwt && exact32)   k2::block_kernel<true, true, false><<<grid, k2::BLOCK_THREADS, 0, s>>>(a);
    else if (!ebwt && !exact32) k2::block_kernel<false, false, false><<<grid, k2::BLOCK_THREADS, 0, s>>>(a);
    else                        k2::block_kernel<false, true, false><<<grid, k2::BLOCK_THREADS, 0, s>>>(a);
    return cudaGetLastError();
}

cudaError_t launch_k2_narrow(const uint32_t *sim32, uint32_t lo, uint32_t hi, uint32_t n_refs,
                             uint8_t *sim8, uint64_t pitch8, cudaStream_t s)
{
    if (hi <= lo || n_refs == 0) return cudaSuccess;
    k2::narrow_kernel<<<148 * 8, 256, 0, s>>>(sim32, lo, hi, n_refs, sim8, pitch8);
    return cudaGetLastError();
}

} // namespace lime
