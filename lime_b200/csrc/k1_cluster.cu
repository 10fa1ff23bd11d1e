// k1_cluster.cu -- step 1, alpha-cluster detection (replaces ClusterLCP.cpp:139-283).
//
// One streaming pass over lcp[] and da[].  A position i with lcp[i] < alpha is a "head"; a
// segment is [head, next head) (the array end closes the last one, ClusterLCP.cpp:244-245).  A
// segment is an alpha-cluster iff it has >= 2 positions (some lcp >= alpha follows the head,
// :216-217), at least one da < n_reads and at least one da >= n_reads (:20-30, :220).
// Positions before the first head of the whole array belong to no segment (:196-202).
//
// Data-parallel form
//   * a CTA takes one tile of 4096 positions (ticket order) and owns the segments whose HEAD is
//     in the tile -- the same ownership rule the reference uses for thread chunks (:150-202) and
//     the one shards use across GPUs;
//   * lcp/da are read with 128-bit loads; each thread turns 4 positions into a 4-bit head nibble
//     (lcp < alpha) and a 4-bit read nibble (da < n_reads); nibbles are paired with one shuffle
//     and stored as bytes, giving two position-ordered bitmaps of the tile (+128 look-ahead
//     positions) in shared memory: 1 KB instead of 33 KB of values;
//   * every thread then walks the heads of its 16 positions with ffs/popc-style bit scans:
//     next head -> length, masked read bitmap -> has-read / has-genome;
//   * a segment still open after the look-ahead window (rare: longer than 128 positions past
//     the tile end) is finished by the whole CTA scanning forward 1024 positions per step
//     (the "run past the chunk end" of ClusterLCP.cpp:246-264);
//   * cluster counts go through a block scan and a decoupled look-back over tiles, so records
//     are written once, compacted, in ascending pStart order.
//
// Algorithmic traffic: 8 B per position read + 16 B per emitted cluster written.
#include "common.cuh"

namespace lime {
namespace k1 {

constexpr int THREADS = 256;
constexpr int TILE = 4096;
constexpr int HALO = 128;
constexpr int EXT = TILE + HALO;     // positions held as bitmaps
constexpr int WORDS = EXT / 32;      // 132
constexpr int GROUPS = EXT / 4;      // 1056 groups of 4 positions
constexpr int LOADS = (GROUPS + THREADS - 1) / THREADS;   // 5
constexpr int NONE = EXT;

__device__ __forceinline__ int next_head(const uint32_t *sH, int i)
{
    int w = i >> 5;
    uint32_t m = sH[w] & ~((2u << (i & 31)) - 1u);
    while (m == 0) {
        if (++w >= WORDS) return NONE;
        m = sH[w];
    }
    return (w << 5) + __ffs(m) - 1;
}

// any read bit / any genome bit among positions [i, j), j > i, j <= EXT
__device__ __forceinline__ void range_rg(const uint32_t *sR, int i, int j, bool &has_r, bool &has_g)
{
    const int w0 = i >> 5, w1 = (j - 1) >> 5;
    const uint32_t m0 = 0xffffffffu << (i & 31);
    const uint32_t m1 = 0xffffffffu >> (31 - ((j - 1) & 31));
    uint32_t r = 0, g = 0;
    if (w0 == w1) {
        const uint32_t m = m0 & m1, b = sR[w0];
        r = b & m; g = ~b & m;
    } else {
        uint32_t b = sR[w0];
        r = b & m0; g = ~b & m0;
        for (int w = w0 + 1; w < w1; w++) { b = sR[w]; r |= b; g |= ~b; }
        b = sR[w1];
        r |= b & m1; g |= ~b & m1;
    }
    has_r = r != 0; has_g = g != 0;
}

__global__ void __launch_bounds__(THREADS) detect_kernel(K1Args P)
{
    __shared__ uint32_t sH[WORDS];
    __shared__ uint32_t sR[WORDS];
    __shared__ uint32_t s_warp_sum[THREADS / 32];
    __shared__ unsigned s_tile;
    __shared__ int s_slow_head;          // tile-local index of the head whose segment leaves the window
    __shared__ unsigned s_min;
    __shared__ unsigned long long s_end; // buffer-local end of that segment
    __shared__ unsigned s_slow_r, s_slow_g;
    __shared__ unsigned long long s_excl;

    const int tid = threadIdx.x;
    if (tid == 0) { s_tile = atomicAdd(P.tile_counter, 1u); s_slow_head = -1; }
    __syncthreads();
    const unsigned tile = s_tile;
    const uint64_t a = (uint64_t)tile * TILE;                  // buffer-local start of the tile
    const int Lt = (int)min((uint64_t)TILE, P.n_own - a);      // owned positions in this tile
    const int Lext = (int)min((uint64_t)EXT, P.n_total - a);   // valid positions in the window

    // ---- load 4224 positions, build the bitmaps -------------------------------------------
    uint4 L[LOADS], D[LOADS];
#pragma unroll
    for (int k = 0; k < LOADS; k++) {
        const int q = tid + k * THREADS;
        if (q < GROUPS) {
            const uint64_t p = a + 4ull * q;
            if (p + 4 <= P.n_total) {
                L[k] = ldg128(P.lcp + p);
                D[k] = ldg128(P.da + p);
            } else {
                uint32_t l[4], d[4];
#pragma unroll
                for (int e = 0; e < 4; e++) {
                    const bool ok = p + e < P.n_total;
                    l[e] = ok ? __ldg(P.lcp + p + e) : 0xffffffffu;   // never a head
                    d[e] = ok ? __ldg(P.da + p + e) : 0u;
                }
                L[k] = make_uint4(l[0], l[1], l[2], l[3]);
                D[k] = make_uint4(d[0], d[1], d[2], d[3]);
            }
        }
    }
    uint8_t *hB = reinterpret_cast<uint8_t *>(sH);
    uint8_t *rB = reinterpret_cast<uint8_t *>(sR);
#pragma unroll
    for (int k = 0; k < LOADS; k++) {
        const int q = tid + k * THREADS;
        if (q < GROUPS) {   // whole warps take this branch together (GROUPS is a multiple of 32)
            unsigned hn = (L[k].x < P.alpha) | ((L[k].y < P.alpha) << 1) | ((L[k].z < P.alpha) << 2) |
                          ((L[k].w < P.alpha) << 3);
            unsigned rn = (D[k].x < P.n_reads) | ((D[k].y < P.n_reads) << 1) |
                          ((D[k].z < P.n_reads) << 2) | ((D[k].w < P.n_reads) << 3);
            const unsigned v = hn | (rn << 4);
            const unsigned o = __shfl_xor_sync(FULL, v, 1);
            if ((q & 1) == 0) hB[q >> 1] = (uint8_t)((v & 0xf) | ((o & 0xf) << 4));
            else              rB[q >> 1] = (uint8_t)((o >> 4) | (v & 0xf0));
        }
    }
    __syncthreads();

    // ---- walk the heads of my 16 positions -------------------------------------------------
    const int base = tid * 16;
    uint32_t heads = 0;
    if (base < Lt) {
        heads = (sH[base >> 5] >> (base & 31)) & 0xffffu;
        if (base + 16 > Lt) heads &= (1u << (Lt - base)) - 1u;
    }
    // the window ends at the end of the array: that end closes the last segment (last shard),
    // or the halo was too short (inner shard)
    const bool window_hits_end = Lext < EXT;

    unsigned count = 0;
    uint64_t max_len = 0;
    bool mine_slow = false, slow_r = false, slow_g = false;
    {
        uint32_t hb = heads;
        while (hb) {
            const int i = base + __ffs(hb) - 1;
            hb &= hb - 1;
            int j = next_head(sH, i);
            if (j == NONE) {
                if (window_hits_end) {
                    if (!P.is_last) atomicOr(P.err, 1u);
                    j = Lext;
                } else {
                    mine_slow = true;
                    range_rg(sR, i, EXT, slow_r, slow_g);
                    s_slow_head = i;
                    continue;
                }
            }
            if (j - i >= 2) {
                bool r, g;
                range_rg(sR, i, j, r, g);
                if (r && g) { count++; max_len = max(max_len, (uint64_t)(j - i)); }
            }
        }
    }
    __syncthreads();

    // ---- rare: finish the segment that runs past the window ---------------------------------
    if (s_slow_head >= 0) {
        uint64_t p0 = a + EXT;
        bool acc_r = false, acc_g = false;
        unsigned long long end = 0;
        while (true) {
            if (tid == 0) s_min = 0xffffffffu;
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
            if (hn) atomicMin(&s_min, 4u * tid + __ffs(hn) - 1);
            __syncthreads();
            const unsigned m = s_min;
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
            if (p0 >= P.n_total) {
                if (!P.is_last && tid == 0) atomicOr(P.err, 1u);
                end = P.n_total;
                break;
            }
        }
        if (tid == 0) { s_end = end; s_slow_r = acc_r; s_slow_g = acc_g; }
        __syncthreads();
        if (mine_slow) {
            const uint64_t len = s_end - (a + s_slow_head);
            if ((slow_r || s_slow_r) && (slow_g || s_slow_g)) { count++; max_len = max(max_len, len); }
            else mine_slow = false;     // not mixed: nothing to emit later
        }
    }

    // ---- block exclusive scan of counts, look-back for the tile offset ----------------------
    const unsigned lane = lane_id(), warp = tid >> 5;
    unsigned incl = count;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const unsigned n = __shfl_up_sync(FULL, incl, o);
        if (lane >= (unsigned)o) incl += n;
    }
    if (lane == 31) s_warp_sum[warp] = incl;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) max_len = max(max_len, __shfl_xor_sync(FULL, max_len, o));
    if (lane == 0 && max_len) atomicMax(P.max_len, (unsigned long long)max_len);
    __syncthreads();
    unsigned warp_off = 0, tile_total = 0;
#pragma unroll
    for (int w = 0; w < THREADS / 32; w++) {
        const unsigned s = s_warp_sum[w];
        if (w < (int)warp) warp_off += s;
        tile_total += s;
    }
    if (warp == 0) {
        const unsigned long long e = lookback_exclusive(P.tile_state, tile, tile_total);
        if (lane == 0) {
            s_excl = e;
            if (tile == gridDim.x - 1) *P.n_clusters = e + tile_total;   // tickets: last tile id
        }
    }
    __syncthreads();
    if (count == 0) return;
    uint64_t o = s_excl + warp_off + (incl - count);

    // ---- emit: same walk, now writing records ------------------------------------------------
    uint32_t hb = heads;
    while (hb) {
        const int i = base + __ffs(hb) - 1;
        hb &= hb - 1;
        int j = next_head(sH, i);
        uint64_t len;
        if (j == NONE && !window_hits_end) {
            if (!mine_slow) continue;
            len = s_end - (a + i);
        } else {
            if (j == NONE) j = Lext;
            if (j - i < 2) continue;
            bool r, g;
            range_rg(sR, i, j, r, g);
            if (!(r && g)) continue;
            len = (uint64_t)(j - i);
        }
        if (o < P.out_cap) {
            ulonglong2 rec = make_ulonglong2(P.pos_base + a + i, len);
            *reinterpret_cast<ulonglong2 *>(&P.out[o]) = rec;
        }
        o++;
    }
}

} // namespace k1

uint64_t k1_num_tiles(uint64_t n_own) { return (n_own + k1::TILE - 1) / k1::TILE; }

cudaError_t launch_k1(const K1Args &a, cudaStream_t s, unsigned *grid_out)
{
    const uint64_t tiles = k1_num_tiles(a.n_own);
    if (grid_out) *grid_out = (unsigned)tiles;
    if (tiles == 0) return cudaSuccess;
    k1::detect_kernel<<<(unsigned)tiles, k1::THREADS, 0, s>>>(a);
    return cudaGetLastError();
}

} // namespace lime
