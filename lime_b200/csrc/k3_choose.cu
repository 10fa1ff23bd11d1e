// k3_choose.cu -- step 2b, per-read threshold and sparse-row emit
// (replaces clusterChoose, ClusterBWT_DA.cpp:361-450, BIN=1).
//
// Row r of the uint8 similarity matrix -> maxSim and the non-zero cells (:390-402);
// normSim = (float)maxSim / norm (:404); if normSim > beta the row is written to .res.bin as a
// header {normSim, count} followed by count x {(float)sim / norm, genome} in ascending genome
// order (:408-426) and .res.pos[r] is the record index of the header, else 0 (:428-434).
// Record 0 is the {0.0f, 0} sentinel (:376-382).
//
// count_kernel : a warp per row (coalesced word loads, __vmaxu4 / byte-nonzero popc), a block
//                scan over 256 reads and a decoupled look-back over tiles give every passing
//                read its record offset in one pass over the matrix; writes pos[].
// emit_kernel  : re-reads the rows (L2 resident for typical sizes) and writes the records.
#include "common.cuh"
#include <algorithm>

namespace lime {
namespace k3 {

constexpr int THREADS = 256;
constexpr int WARPS = THREADS / 32;
constexpr int TILE_READS = 256;               // reads per CTA
constexpr int ROWS_PER_WARP = TILE_READS / WARPS;

// per-byte "non-zero" mask: 0x01 in every byte of w that is non-zero
__device__ __forceinline__ uint32_t nz_bytes(uint32_t w)
{
    return __vcmpne4(w, 0u) & 0x01010101u;
}

__global__ void __launch_bounds__(THREADS) count_kernel(K3Args P)
{
    __shared__ uint32_t s_cnt[TILE_READS];
    __shared__ uint32_t s_warp_sum[WARPS];
    __shared__ unsigned s_tile;
    __shared__ unsigned long long s_excl;
    const int tid = threadIdx.x;
    const unsigned lane = lane_id(), warp = tid >> 5;
    if (tid == 0) s_tile = atomicAdd(P.tile_counter, 1u);
    __syncthreads();
    const unsigned tile = s_tile;
    const uint32_t n = P.read_end - P.read_begin;
    const uint32_t r0 = tile * TILE_READS;
    const uint32_t words = (P.n_refs + 3) / 4;
    const float normf = (float)P.norm;

    for (int k = 0; k < ROWS_PER_WARP; k++) {
        const uint32_t lr = r0 + warp * ROWS_PER_WARP + k;
        uint32_t mx = 0, nnz = 0;
        if (lr < n) {
            const uint32_t *row = reinterpret_cast<const uint32_t *>(
                P.sim8 + (size_t)(P.read_begin + lr) * P.pitch8);
            for (uint32_t w = lane; w < words; w += 32) {
                const uint32_t v = __ldg(row + w);
                mx = __vmaxu4(mx, v);
                nnz += __popc(nz_bytes(v));
            }
        }
        mx = max(max(mx & 0xff, (mx >> 8) & 0xff), max((mx >> 16) & 0xff, mx >> 24));
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            mx = max(mx, __shfl_xor_sync(FULL, mx, o));
            nnz += __shfl_xor_sync(FULL, nnz, o);
        }
        if (lane == 0) {
            const float ns = __fdiv_rn((float)mx, normf);
            s_cnt[warp * ROWS_PER_WARP + k] = (lr < n && ns > P.beta) ? 1u + nnz : 0u;
        }
    }
    __syncthreads();
    // block exclusive scan over the 256 per-read record counts
    const uint32_t c = s_cnt[tid];
    uint32_t incl = c;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t v = __shfl_up_sync(FULL, incl, o);
        if (lane >= (unsigned)o) incl += v;
    }
    if (lane == 31) s_warp_sum[warp] = incl;
    __syncthreads();
    uint32_t warp_off = 0, total = 0;
#pragma unroll
    for (int w = 0; w < WARPS; w++) {
        const uint32_t s = s_warp_sum[w];
        if (w < (int)warp) warp_off += s;
        total += s;
    }
    if (warp == 0) {
        const unsigned long long e = lookback_exclusive(P.tile_state, tile, total);
        if (lane == 0) {
            s_excl = e;
            if (tile == gridDim.x - 1) *P.n_records = 1ull + e + total;
        }
    }
    __syncthreads();
    const uint32_t lr = r0 + tid;
    if (lr < n) P.pos[lr] = c ? 1ull + s_excl + warp_off + (incl - c) : 0ull;
}

__global__ void __launch_bounds__(THREADS) emit_kernel(K3Args P)
{
    const unsigned lane = lane_id();
    const uint32_t n = P.read_end - P.read_begin;
    const uint64_t warp_global = ((uint64_t)blockIdx.x * THREADS + threadIdx.x) >> 5;
    const uint64_t warp_stride = ((uint64_t)gridDim.x * THREADS) >> 5;
    const uint32_t words = (P.n_refs + 3) / 4;
    const float normf = (float)P.norm;
    if (warp_global == 0 && lane == 0) { P.bin[0].sim = 0.0f; P.bin[0].idRef = 0; }

    for (uint64_t lr = warp_global; lr < n; lr += warp_stride) {
        const uint64_t o = P.pos[lr];
        if (o == 0) continue;
        const uint32_t *row = reinterpret_cast<const uint32_t *>(
            P.sim8 + (size_t)(P.read_begin + lr) * P.pitch8);
        uint32_t mx = 0, written = 0;
        for (uint32_t w0 = 0; w0 < words; w0 += 32) {
            const uint32_t w = w0 + lane;
            const uint32_t v = w < words ? __ldg(row + w) : 0u;
            mx = __vmaxu4(mx, v);
            const uint32_t cnt = __popc(nz_bytes(v));
            uint32_t incl = cnt;
#pragma unroll
            for (int s = 1; s < 32; s <<= 1) {
                const uint32_t x = __shfl_up_sync(FULL, incl, s);
                if (lane >= (unsigned)s) incl += x;
            }
            uint64_t dst = o + 1 + written + (incl - cnt);
#pragma unroll
            for (int b = 0; b < 4; b++) {
                const uint32_t cell = (v >> (8 * b)) & 0xffu;
                if (cell) {
                    lime_pair_sim rec;
                    rec.sim = __fdiv_rn((float)cell, normf);
                    rec.idRef = 4 * w + b;
                    P.bin[dst++] = rec;
                }
            }
            written += __shfl_sync(FULL, incl, 31);
        }
        mx = max(max(mx & 0xff, (mx >> 8) & 0xff), max((mx >> 16) & 0xff, mx >> 24));
#pragma unroll
        for (int s = 16; s > 0; s >>= 1) mx = max(mx, __shfl_xor_sync(FULL, mx, s));
        if (lane == 0) {
            lime_pair_sim h;
            h.sim = __fdiv_rn((float)mx, normf);
            h.idRef = written;
            P.bin[o] = h;
        }
    }
}


__global__ void import_kernel(const lime_pair_sim *tr, uint64_t n, uint8_t *sim8, uint64_t pitch8, uint32_t n_reads,
                              uint32_t n_refs, unsigned int *err)
{
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x) {
        const uint2 t = __ldg(reinterpret_cast<const uint2 *>(tr) + i);
        const uint32_t r = t.x, g = t.y >> 8;
        if (r >= n_reads || g >= n_refs) { atomicOr(err, 1u); continue; }
        uint8_t *cell = sim8 + (size_t)r * pitch8 + g;
        *cell = (uint8_t)(*cell + (t.y & 0xffu));
    }
}

// ---- single pass: count, offsets (look-back) and emit in one kernel ---------------------------------
// Used when an upper bound of the record count is known beforehand (1 + reads + similarity updates),
// so the record buffer can be sized without a counting pass.  A row is read once from HBM; the emit
// step re-reads the CTA's 256 rows from L1/L2.  LPR lanes share a row (16 bytes per lane and step).
__device__ __forceinline__ uint32_t sum_bytes(const uint4 &v, uint32_t acc)
{
    acc = __dp4a(v.x, 0x01010101u, acc); acc = __dp4a(v.y, 0x01010101u, acc);
    acc = __dp4a(v.z, 0x01010101u, acc); acc = __dp4a(v.w, 0x01010101u, acc);
    return acc;
}
// one atomic per CTA for the byte sum of the rows it read (carry proof of the fire-and-forget adds)
__device__ __forceinline__ void publish_byte_sum(unsigned long long *out, uint32_t mine, unsigned long long *s_part)
{
    unsigned long long v = mine;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(FULL, v, o);
    if ((threadIdx.x & 31) == 0) s_part[threadIdx.x >> 5] = v;
    __syncthreads();
    if (threadIdx.x == 0) {
        unsigned long long t = 0;
        for (int k = 0; k < WARPS; k++) t += s_part[k];
        if (t) atomicAdd(out, t);
    }
}

template <int LPR>
__device__ __forceinline__ void row_stats(const uint8_t *row, uint32_t chunks, unsigned sub, bool on,
                                          uint32_t &mx, uint32_t &nnz, uint32_t &bsum)
{
    uint32_t m = 0;
    nnz = 0;
    if (on) {
        for (uint32_t c = sub; c < chunks; c += LPR) {
            const uint4 v = __ldg(reinterpret_cast<const uint4 *>(row) + c);
            m = __vmaxu4(__vmaxu4(m, v.x), __vmaxu4(__vmaxu4(v.y, v.z), v.w));
            nnz += __popc(nz_bytes(v.x)) + __popc(nz_bytes(v.y)) + __popc(nz_bytes(v.z)) + __popc(nz_bytes(v.w));
            bsum = sum_bytes(v, bsum);
        }
    }
    __syncwarp();
    mx = max(max(m & 0xff, (m >> 8) & 0xff), max((m >> 16) & 0xff, m >> 24));
#pragma unroll
    for (int o = LPR / 2; o > 0; o >>= 1) {
        mx = max(mx, __shfl_xor_sync(FULL, mx, o));
        nnz += __shfl_xor_sync(FULL, nnz, o);
    }
}

// TRIPLES: instead of .res records, export every non-zero cell of the rows as a triple
// {read id, genome << 8 | count} (no header, no beta test) -- the multi-GPU exchange format.
template <int LPR>
__global__ void __launch_bounds__(THREADS) choose_kernel(K3Args P)
{
    constexpr bool TRIPLES = false;      // (triples are exported by export_kernel)
    constexpr unsigned HDR = TRIPLES ? 0u : 1u;
    constexpr int RPW = 32 / LPR;                    // rows a warp handles per step
    __shared__ uint32_t s_cnt[TILE_READS];
    __shared__ uint32_t s_off[TILE_READS];
    __shared__ uint32_t s_warp_sum[WARPS];
    __shared__ unsigned long long s_part[WARPS];
    __shared__ unsigned s_tile;
    __shared__ unsigned long long s_excl;
    const int tid = threadIdx.x;
    const unsigned lane = lane_id(), warp = tid >> 5;
    const unsigned sub = lane % LPR, grp = lane / LPR;
    if (tid == 0) s_tile = atomicAdd(P.tile_counter, 1u);
    __syncthreads();
    const unsigned tile = s_tile;
    const uint32_t n = P.read_end - P.read_begin;
    const uint32_t r0 = tile * TILE_READS;
    const uint32_t chunks = (uint32_t)(P.pitch8 / 16);
    const float normf = (float)P.norm;
    if (!TRIPLES && tile == 0 && tid == 0 && P.bin_cap) { P.bin[0].sim = 0.0f; P.bin[0].idRef = 0; }   // sentinel (:376-382)
    uint32_t bsum = 0;                                // byte sum of the cells this thread read (< 2^32: <= 2 MB per thread)

    // ---- count ----
    for (int k = 0; k < ROWS_PER_WARP; k += RPW) {
        const uint32_t lr = r0 + warp * ROWS_PER_WARP + k + grp;
        uint32_t mx = 0, nnz = 0;
        row_stats<LPR>(P.sim8 + (size_t)(P.read_begin + lr) * P.pitch8, chunks, sub, lr < n, mx, nnz, bsum);
        if (sub == 0) {
            const float ns = __fdiv_rn((float)mx, normf);
            const bool pass = TRIPLES ? (nnz != 0) : (ns > P.beta);
            s_cnt[warp * ROWS_PER_WARP + k + grp] = (lr < n && pass) ? HDR + nnz : 0u;
        }
    }
    __syncthreads();
    // ---- offsets ----
    const uint32_t c = s_cnt[tid];
    uint32_t incl = c;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t v = __shfl_up_sync(FULL, incl, o);
        if (lane >= (unsigned)o) incl += v;
    }
    if (lane == 31) s_warp_sum[warp] = incl;
    __syncthreads();
    uint32_t warp_off = 0, total = 0;
#pragma unroll
    for (int w = 0; w < WARPS; w++) {
        const uint32_t s = s_warp_sum[w];
        if (w < (int)warp) warp_off += s;
        total += s;
    }
    if (warp == 0) {
        const unsigned long long e = lookback_exclusive(P.tile_state, tile, total);
        if (lane == 0) {
            s_excl = e;
            if (tile == gridDim.x - 1) *P.n_records = HDR + e + total;
        }
    }
    s_off[tid] = warp_off + (incl - c);
    __syncthreads();
    const unsigned long long tile_base = HDR + s_excl;
    if (!TRIPLES) {
        const uint32_t lr = r0 + tid;
        if (lr < n) P.pos[lr] = c ? tile_base + s_off[tid] : 0ull;
    }
    // ---- emit ----
    for (int k = 0; k < ROWS_PER_WARP; k += RPW) {
        const int slot = warp * ROWS_PER_WARP + k + grp;
        const uint32_t lr = r0 + slot;
        const uint32_t cnt = s_cnt[slot];
        const bool on = lr < n && cnt != 0;
        uint8_t *row = P.sim8 + (size_t)(P.read_begin + lr) * P.pitch8;
        const unsigned long long o = tile_base + s_off[slot];
        uint32_t mx = 0, written = 0;
        if (!P.zero_after && !__any_sync(FULL, on)) continue;
        for (uint32_t c0 = 0; c0 < chunks; c0 += LPR) {
            const uint32_t ch = c0 + sub;
            uint4 v = make_uint4(0, 0, 0, 0);
            if ((on || (P.zero_after && lr < n)) && ch < chunks) {
                v = *(reinterpret_cast<const uint4 *>(row) + ch);
                // leave the matrix zero for the next strand (only the 16-byte pieces that hold something)
                if (P.zero_after && (v.x | v.y | v.z | v.w)) *(reinterpret_cast<uint4 *>(row) + ch) = make_uint4(0, 0, 0, 0);
                if (!on) v = make_uint4(0, 0, 0, 0);
            }
            const uint32_t wv[4] = {v.x, v.y, v.z, v.w};
            uint32_t mine = 0;
#pragma unroll
            for (int x = 0; x < 4; x++) { mx = __vmaxu4(mx, wv[x]); mine += __popc(nz_bytes(wv[x])); }
            uint32_t inc2 = mine;
#pragma unroll
            for (int s = 1; s < LPR; s <<= 1) {
                const uint32_t y = __shfl_up_sync(FULL, inc2, s, LPR);
                if (sub >= (unsigned)s) inc2 += y;
            }
            unsigned long long dst = o + HDR + written + (inc2 - mine);
            if (mine) {
#pragma unroll
                for (int x = 0; x < 4; x++) {
#pragma unroll
                    for (int b = 0; b < 4; b++) {
                        const uint32_t cell = (wv[x] >> (8 * b)) & 0xffu;
                        if (cell) {
                            if (dst < P.bin_cap) {
                                lime_pair_sim rec;
                                if (TRIPLES) {
                                    rec.sim = __uint_as_float(P.read_begin + lr);
                                    rec.idRef = ((16 * ch + 4 * x + b) << 8) | cell;
                                } else {
                                    rec.sim = __fdiv_rn((float)cell, normf);
                                    rec.idRef = 16 * ch + 4 * x + b;
                                }
                                P.bin[dst] = rec;
                            }
                            dst++;
                        }
                    }
                }
            }
            written += __shfl_sync(FULL, inc2, LPR - 1, LPR);
        }
        mx = max(max(mx & 0xff, (mx >> 8) & 0xff), max((mx >> 16) & 0xff, mx >> 24));
#pragma unroll
        for (int s = LPR / 2; s > 0; s >>= 1) mx = max(mx, __shfl_xor_sync(FULL, mx, s));
        if (!TRIPLES && on && sub == 0 && o < P.bin_cap) {
            lime_pair_sim h;
            h.sim = __fdiv_rn((float)mx, normf);
            h.idRef = written;
            P.bin[o] = h;
        }
    }
    if (P.byte_sum) publish_byte_sum(P.byte_sum, bsum, s_part);
}

// ---- single pass for rows of at most 128 bytes (n_refs <= 128): sparse-aware ----------------------
// The CTA's 256 rows are loaded once (all loads of a thread in flight together) into shared memory;
// a lane turns its 16 cells into a 16-bit non-zero mask and only walks the set bits, so the work is
// proportional to the non-zero cells (a few per row) instead of n_refs.
__device__ __forceinline__ uint32_t nz_nibble(uint32_t w)      // bit b: byte b of w is non-zero
{
    const uint32_t h = (((w & 0x7f7f7f7fu) + 0x7f7f7f7fu) | w) & 0x80808080u;
    return ((h >> 7) * 0x01020408u) >> 24;
}
__device__ __forceinline__ uint32_t nz_mask16(const uint4 &v)
{
    return nz_nibble(v.x) | (nz_nibble(v.y) << 4) | (nz_nibble(v.z) << 8) | (nz_nibble(v.w) << 12);
}

template <int LPR>
__global__ void __launch_bounds__(THREADS) choose_sparse_kernel(K3Args P)
{
    constexpr bool TRIPLES = false;      // (triples are exported by export_kernel)
    constexpr unsigned HDR = TRIPLES ? 0u : 1u;
    constexpr int RPW = 32 / LPR;                    // rows a warp handles per step
    constexpr int STEPS = ROWS_PER_WARP / RPW;
    __shared__ uint4 s_rows[TILE_READS * LPR];
    __shared__ uint32_t s_cnt[TILE_READS];
    __shared__ uint32_t s_off[TILE_READS];
    __shared__ uint32_t s_max[TILE_READS];
    __shared__ uint32_t s_warp_sum[WARPS];
    __shared__ unsigned long long s_part[WARPS];
    __shared__ unsigned s_tile;
    __shared__ unsigned long long s_excl;
    const int tid = threadIdx.x;
    const unsigned lane = lane_id(), warp = tid >> 5;
    const unsigned sub = lane % LPR, grp = lane / LPR;
    if (tid == 0) s_tile = atomicAdd(P.tile_counter, 1u);
    __syncthreads();
    const unsigned tile = s_tile;
    const uint32_t n = P.read_end - P.read_begin;
    const uint32_t r0 = tile * TILE_READS;
    const uint32_t chunks = (uint32_t)(P.pitch8 / 16);           // <= LPR
    const float normf = (float)P.norm;
    const uint8_t *s_bytes = reinterpret_cast<const uint8_t *>(s_rows);
    if (!TRIPLES && tile == 0 && tid == 0 && P.bin_cap) { P.bin[0].sim = 0.0f; P.bin[0].idRef = 0; }   // sentinel (:376-382)

    // ---- load ----
    {
        uint4 held[STEPS];
#pragma unroll
        for (int st = 0; st < STEPS; st++) {
            const uint32_t lr = r0 + warp * ROWS_PER_WARP + st * RPW + grp;
            held[st] = make_uint4(0, 0, 0, 0);
            if (lr < n && sub < chunks)
                held[st] = __ldcs(reinterpret_cast<const uint4 *>(P.sim8 + (size_t)(P.read_begin + lr) * P.pitch8) + sub);
        }
        uint32_t bsum = 0;
#pragma unroll
        for (int st = 0; st < STEPS; st++) {
            s_rows[(warp * ROWS_PER_WARP + st * RPW + grp) * LPR + sub] = held[st];
            bsum = sum_bytes(held[st], bsum);
            // leave the matrix zero for the next strand (only the 16-byte pieces that hold something)
            if (P.zero_after && (held[st].x | held[st].y | held[st].z | held[st].w)) {
                const uint32_t lr = r0 + warp * ROWS_PER_WARP + st * RPW + grp;
                *(reinterpret_cast<uint4 *>(P.sim8 + (size_t)(P.read_begin + lr) * P.pitch8) + sub) = make_uint4(0, 0, 0, 0);
            }
        }
        if (P.byte_sum) publish_byte_sum(P.byte_sum, bsum, s_part);     // every thread of the CTA gets here
        __syncwarp();
    }
    // ---- count ----
    for (int k = 0; k < ROWS_PER_WARP; k += RPW) {
        const int slot = warp * ROWS_PER_WARP + k + grp;
        const uint32_t lr = r0 + slot;
        uint32_t m16 = nz_mask16(s_rows[slot * LPR + sub]);
        uint32_t nnz = __popc(m16), mx = 0;
        while (m16) {
            const int i = __ffs(m16) - 1;
            m16 &= m16 - 1;
            mx = max(mx, (uint32_t)s_bytes[(slot * LPR + sub) * 16 + i]);
        }
#pragma unroll
        for (int o = LPR / 2; o > 0; o >>= 1) {
            mx = max(mx, __shfl_xor_sync(FULL, mx, o));
            nnz += __shfl_xor_sync(FULL, nnz, o);
        }
        if (sub == 0) {
            const float ns = __fdiv_rn((float)mx, normf);
            const bool pass = TRIPLES ? (nnz != 0) : (ns > P.beta);
            s_cnt[slot] = (lr < n && pass) ? HDR + nnz : 0u;
            s_max[slot] = mx;
        }
    }
    __syncthreads();
    // ---- offsets ----
    const uint32_t c = s_cnt[tid];
    uint32_t incl = c;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t v = __shfl_up_sync(FULL, incl, o);
        if (lane >= (unsigned)o) incl += v;
    }
    if (lane == 31) s_warp_sum[warp] = incl;
    __syncthreads();
    uint32_t warp_off = 0, total = 0;
#pragma unroll
    for (int w = 0; w < WARPS; w++) {
        const uint32_t x = s_warp_sum[w];
        if (w < (int)warp) warp_off += x;
        total += x;
    }
    if (warp == 0) {
        const unsigned long long e = lookback_exclusive(P.tile_state, tile, total);
        if (lane == 0) {
            s_excl = e;
            if (tile == gridDim.x - 1) *P.n_records = HDR + e + total;
        }
    }
    s_off[tid] = warp_off + (incl - c);
    __syncthreads();
    const unsigned long long tile_base = HDR + s_excl;
    if (!TRIPLES) {
        const uint32_t lr = r0 + tid;
        if (lr < n) P.pos[lr] = c ? tile_base + s_off[tid] : 0ull;
    }
    // ---- emit ----
    for (int k = 0; k < ROWS_PER_WARP; k += RPW) {
        const int slot = warp * ROWS_PER_WARP + k + grp;
        const uint32_t lr = r0 + slot;
        const bool on = s_cnt[slot] != 0;                       // implies lr < n
        if (!__any_sync(FULL, on)) continue;
        const unsigned long long o = tile_base + s_off[slot];
        uint32_t m16 = on ? nz_mask16(s_rows[slot * LPR + sub]) : 0u;
        const uint32_t mine = __popc(m16);
        uint32_t inc2 = mine;
#pragma unroll
        for (int x = 1; x < LPR; x <<= 1) {
            const uint32_t y = __shfl_up_sync(FULL, inc2, x, LPR);
            if (sub >= (unsigned)x) inc2 += y;
        }
        unsigned long long dst = o + HDR + (inc2 - mine);
        while (m16) {
            const int i = __ffs(m16) - 1;
            m16 &= m16 - 1;
            const uint32_t cell = s_bytes[(slot * LPR + sub) * 16 + i];
            if (dst < P.bin_cap) {
                lime_pair_sim rec;
                if (TRIPLES) {
                    rec.sim = __uint_as_float(P.read_begin + lr);
                    rec.idRef = ((16 * sub + i) << 8) | cell;
                } else {
                    rec.sim = __fdiv_rn((float)cell, normf);
                    rec.idRef = 16 * sub + i;
                }
                P.bin[dst] = rec;
            }
            dst++;
        }
        if (!TRIPLES && on && sub == 0 && o < P.bin_cap) {
            lime_pair_sim h;
            h.sim = __fdiv_rn((float)s_max[slot], normf);
            h.idRef = s_cnt[slot] - HDR;
            P.bin[o] = h;
        }
    }
}

// ---- single pass for rows of 129 .. 1024 genomes (9 .. 64 sixteen-byte pieces): sparse-aware, streaming ---------
// The shapes of BASELINE configs 3 and 4 (930 / 1000 genomes): a row is 1 KB but holds a handful of non-zero
// cells.  Phase 1 streams the CTA's 256 rows once -- four rows (up to eight 16-byte loads per lane) in flight
// per warp -- and keeps only the non-zero pieces, in (row, piece) order, in a per-warp shared-memory stash; a
// piece is cleared in the matrix as soon as it is stashed (zero_after).  Phases 2 and 3 work on the stash with
// one lane per row: row maximum and record count, then (after the block scan + look-back that places the tile)
// the records.  A warp whose stash is full leaves the remaining rows to a plain re-read from global memory.
constexpr int WIDE_STASH = 256;                   // pieces per warp (32 rows)
constexpr unsigned WIDE_REREAD = 0xffffu;

template <int NB>                                 // 32-piece blocks per row: 1 (<= 32 pieces) or 2 (<= 64)
__global__ void __launch_bounds__(THREADS) choose_wide_kernel(K3Args P)
{
    __shared__ uint4 s_val[WARPS][WIDE_STASH];
    __shared__ uint16_t s_piece[WARPS][WIDE_STASH];
    __shared__ uint32_t s_cnt[TILE_READS];
    __shared__ uint32_t s_off[TILE_READS];
    __shared__ uint32_t s_max[TILE_READS];
    __shared__ uint16_t s_first[TILE_READS], s_nent[TILE_READS];
    __shared__ uint32_t s_warp_sum[WARPS];
    __shared__ unsigned long long s_part[WARPS];
    __shared__ unsigned s_tile;
    __shared__ unsigned long long s_excl;
    const int tid = threadIdx.x;
    const unsigned lane = lane_id(), warp = tid >> 5;
    const uint32_t lt_mask = (1u << lane) - 1u;
    if (tid == 0) s_tile = atomicAdd(P.tile_counter, 1u);
    __syncthreads();
    const unsigned tile = s_tile;
    const uint32_t n = P.read_end - P.read_begin;
    const uint32_t r0 = tile * TILE_READS;
    const uint32_t chunks = (uint32_t)(P.pitch8 / 16);           // <= 32 * NB
    const float normf = (float)P.norm;
    if (tile == 0 && tid == 0 && P.bin_cap) { P.bin[0].sim = 0.0f; P.bin[0].idRef = 0; }   // sentinel (:376-382)

    // ---- phase 1: stream the warp's 32 rows, stash the non-zero pieces ----
    uint32_t bsum = 0;
    unsigned cntw = 0;                                           // pieces stashed by this warp (warp-uniform)
    constexpr int RF = 4;                                        // rows in flight
    for (int k0 = 0; k0 < ROWS_PER_WARP; k0 += RF) {
        uint4 v[RF][NB];
#pragma unroll
        for (int j = 0; j < RF; j++) {
            const uint32_t lr = r0 + warp * ROWS_PER_WARP + k0 + j;
            const uint4 *row = reinterpret_cast<const uint4 *>(P.sim8 + (size_t)(P.read_begin + lr) * P.pitch8);
#pragma unroll
            for (int b = 0; b < NB; b++) {
                const uint32_t c = 32u * b + lane;
                v[j][b] = (lr < n && c < chunks) ? __ldcs(row + c) : make_uint4(0, 0, 0, 0);
            }
        }
#pragma unroll
        for (int j = 0; j < RF; j++) {
            const int slot = warp * ROWS_PER_WARP + k0 + j;
            const uint32_t lr = r0 + slot;
            unsigned m[NB], total = 0;
#pragma unroll
            for (int b = 0; b < NB; b++) {
                m[b] = __ballot_sync(FULL, (v[j][b].x | v[j][b].y | v[j][b].z | v[j][b].w) != 0u);
                total += __popc(m[b]);
            }
            const bool fits = cntw + total <= (unsigned)WIDE_STASH;
            if (lane == 0) { s_first[slot] = (uint16_t)cntw; s_nent[slot] = fits ? (uint16_t)total : (uint16_t)WIDE_REREAD; }
            if (!fits) continue;                                  // warp-uniform: this row is re-read in phase 2 / 3
#pragma unroll
            for (int b = 0; b < NB; b++) {
                if ((m[b] >> lane) & 1u) {
                    const unsigned e = cntw + __popc(m[b] & lt_mask);
                    s_val[warp][e] = v[j][b];
                    s_piece[warp][e] = (uint16_t)(32u * b + lane);
                    bsum = sum_bytes(v[j][b], bsum);
                    // leave the matrix zero for the next strand
                    if (P.zero_after)
                        *(reinterpret_cast<uint4 *>(P.sim8 + (size_t)(P.read_begin + lr) * P.pitch8) + 32u * b + lane) = make_uint4(0, 0, 0, 0);
                }
                cntw += __popc(m[b]);
            }
        }
    }
    __syncwarp();
    // ---- phase 2: one lane per row: maximum and number of non-zero cells ----
    const int slot = warp * ROWS_PER_WARP + (int)lane;
    const uint32_t lr = r0 + slot;
    const unsigned first = s_first[slot], nent = s_nent[slot];
    const uint8_t *rowg = P.sim8 + (size_t)(P.read_begin + lr) * P.pitch8;
    {
        uint32_t mx = 0, nnz = 0;
        if (nent != WIDE_REREAD) {
            for (unsigned e = first; e < first + nent; e++) {
                const uint4 x = s_val[warp][e];
                mx = __vmaxu4(__vmaxu4(mx, x.x), __vmaxu4(__vmaxu4(x.y, x.z), x.w));
                nnz += __popc(nz_mask16(x));
            }
        } else if (lr < n) {
            for (uint32_t c = 0; c < chunks; c++) {
                const uint4 x = *(reinterpret_cast<const uint4 *>(rowg) + c);
                mx = __vmaxu4(__vmaxu4(mx, x.x), __vmaxu4(__vmaxu4(x.y, x.z), x.w));
                nnz += __popc(nz_mask16(x));
                bsum = sum_bytes(x, bsum);
            }
        }
        mx = max(max(mx & 0xff, (mx >> 8) & 0xff), max((mx >> 16) & 0xff, mx >> 24));
        const float ns = __fdiv_rn((float)mx, normf);
        s_cnt[slot] = (lr < n && ns > P.beta) ? 1u + nnz : 0u;
        s_max[slot] = mx;
    }
    __syncthreads();
    // ---- offsets ----
    const uint32_t c = s_cnt[tid];
    uint32_t incl = c;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t y = __shfl_up_sync(FULL, incl, o);
        if (lane >= (unsigned)o) incl += y;
    }
    if (lane == 31) s_warp_sum[warp] = incl;
    __syncthreads();
    uint32_t warp_off = 0, total = 0;
#pragma unroll
    for (int w = 0; w < WARPS; w++) {
        const uint32_t x = s_warp_sum[w];
        if (w < (int)warp) warp_off += x;
        total += x;
    }
    if (warp == 0) {
        const unsigned long long e = lookback_exclusive(P.tile_state, tile, total);
        if (lane == 0) {
            s_excl = e;
            if (tile == gridDim.x - 1) *P.n_records = 1ull + e + total;
        }
    }
    s_off[tid] = warp_off + (incl - c);
    __syncthreads();
    const unsigned long long tile_base = 1ull + s_excl;
    {
        const uint32_t lr2 = r0 + tid;
        if (lr2 < n) P.pos[lr2] = c ? tile_base + s_off[tid] : 0ull;
    }
    // ---- phase 3: one lane per row: the records, ascending genome order (pieces are stashed in that order) ----
    {
        const bool on = s_cnt[slot] != 0;                        // implies lr < n
        unsigned long long dst = tile_base + s_off[slot];
        if (on && dst < P.bin_cap) {
            lime_pair_sim h;
            h.sim = __fdiv_rn((float)s_max[slot], normf);
            h.idRef = s_cnt[slot] - 1u;
            P.bin[dst] = h;
        }
        dst++;
        auto emit_piece = [&](const uint4 &x, uint32_t piece) {
            const uint32_t wv[4] = {x.x, x.y, x.z, x.w};
            uint32_t mm = nz_mask16(x);
            while (mm) {
                const int b = __ffs(mm) - 1;
                mm &= mm - 1;
                const uint32_t cell = (wv[b >> 2] >> ((b & 3) * 8)) & 0xffu;
                if (dst < P.bin_cap) {
                    lime_pair_sim rec;
                    rec.sim = __fdiv_rn((float)cell, normf);
                    rec.idRef = 16u * piece + (uint32_t)b;
                    P.bin[dst] = rec;
                }
                dst++;
            }
        };
        if (nent != WIDE_REREAD) {
            if (on)
                for (unsigned e = first; e < first + nent; e++) emit_piece(s_val[warp][e], s_piece[warp][e]);
        } else if (lr < n) {
            for (uint32_t cc = 0; cc < chunks; cc++) {
                uint4 *p = reinterpret_cast<uint4 *>(const_cast<uint8_t *>(rowg)) + cc;
                const uint4 x = *p;
                if ((x.x | x.y | x.z | x.w) == 0u) continue;
                if (on) emit_piece(x, cc);
                if (P.zero_after) *p = make_uint4(0, 0, 0, 0);
            }
        }
    }
    if (P.byte_sum) publish_byte_sum(P.byte_sum, bsum, s_part);
}

// ---- export of the non-zero cells of a row range as triples, in no particular order ----------------
// (the multi-GPU exchange format {read id, genome << 8 | count}; the importer adds them in any order).
// A pure streaming pass: one 16-cell chunk per lane, warp-aggregated append with one atomic per warp.
constexpr int EXPORT_UNROLL = 8;          // chunks per lane and round: one append atomic per 4 KB of matrix
__global__ void __launch_bounds__(THREADS) export_kernel(K3Args P)
{
    const unsigned lane = lane_id();
    const uint64_t chunks = P.pitch8 / 16;
    const uint64_t items = (uint64_t)(P.read_end - P.read_begin) * chunks;
    const uint64_t per_warp = 32ull * EXPORT_UNROLL;
    const uint64_t n_groups = (items + per_warp - 1) / per_warp;
    const uint64_t warp_global = ((uint64_t)blockIdx.x * THREADS + threadIdx.x) >> 5;
    const uint64_t warp_stride = ((uint64_t)gridDim.x * THREADS) >> 5;
    const uint4 *base16 = reinterpret_cast<const uint4 *>(P.sim8 + (size_t)P.read_begin * P.pitch8);   // rows are contiguous
    for (uint64_t grp = warp_global; grp < n_groups; grp += warp_stride) {
        uint4 v[EXPORT_UNROLL];
        uint32_t m[EXPORT_UNROLL];
        uint32_t cnt = 0;
#pragma unroll
        for (int u = 0; u < EXPORT_UNROLL; u++) {
            const uint64_t i = grp * per_warp + 32ull * u + lane;
            v[u] = i < items ? __ldcs(base16 + i) : make_uint4(0, 0, 0, 0);
        }
#pragma unroll
        for (int u = 0; u < EXPORT_UNROLL; u++) { m[u] = nz_mask16(v[u]); cnt += __popc(m[u]); }
        if (!__any_sync(FULL, cnt != 0)) continue;
        uint32_t incl = cnt;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t y = __shfl_up_sync(FULL, incl, o);
            if (lane >= (unsigned)o) incl += y;
        }
        unsigned long long dst = 0;
        if (lane == 31) dst = atomicAdd(P.n_records, (unsigned long long)incl);
        dst = __shfl_sync(FULL, dst, 31) + (incl - cnt);
#pragma unroll
        for (int u = 0; u < EXPORT_UNROLL; u++) {
            if (!m[u]) continue;
            const uint64_t i = grp * per_warp + 32ull * u + lane;
            const uint32_t row = (uint32_t)(i / chunks), ch = (uint32_t)(i % chunks);
            const uint32_t wv[4] = {v[u].x, v[u].y, v[u].z, v[u].w};
            uint32_t mm = m[u];
            while (mm) {
                const int b = __ffs(mm) - 1;
                mm &= mm - 1;
                const uint32_t cell = (wv[b >> 2] >> ((b & 3) * 8)) & 0xffu;
                if (dst < P.bin_cap) {
                    lime_pair_sim rec;
                    rec.sim = __uint_as_float(P.read_begin + row);
                    rec.idRef = ((16 * ch + b) << 8) | cell;
                    P.bin[dst] = rec;
                }
                dst++;
            }
        }
    }
}

// ---- byte sum of the matrix (carry check of the fire-and-forget adds when no choose pass follows) -------
constexpr int SUM_THREADS = 256;
__global__ void __launch_bounds__(SUM_THREADS) bytesum_kernel(const uint4 *p, uint64_t n16, unsigned long long *out)
{
    unsigned acc4 = 0;                                          // <= 2^32 / 1020 vectors per thread
    unsigned long long total = 0;
    for (uint64_t i = (uint64_t)blockIdx.x * SUM_THREADS + threadIdx.x; i < n16; i += (uint64_t)gridDim.x * SUM_THREADS) {
        acc4 = sum_bytes(__ldcs(p + i), acc4);
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

} // namespace k3

// bytes must be a multiple of 16 and p 16-byte aligned (the matrix pitch is)
cudaError_t launch_bytesum(const uint8_t *p, uint64_t bytes, unsigned long long *out, cudaStream_t s)
{
    const uint64_t n16 = bytes / 16;
    if (!n16) return cudaSuccess;
    const unsigned grid = (unsigned)std::min<uint64_t>((n16 + k3::SUM_THREADS - 1) / k3::SUM_THREADS, 148ull * 16);
    k3::bytesum_kernel<<<grid, k3::SUM_THREADS, 0, s>>>(reinterpret_cast<const uint4 *>(p), n16, out);
    return cudaGetLastError();
}

uint64_t k3_num_tiles(uint32_t n_reads) { return ((uint64_t)n_reads + k3::TILE_READS - 1) / k3::TILE_READS; }

cudaError_t launch_k3_count(const K3Args &a, cudaStream_t s)
{
    const uint64_t tiles = k3_num_tiles(a.read_end - a.read_begin);
    if (tiles == 0) return cudaSuccess;
    k3::count_kernel<<<(unsigned)tiles, k3::THREADS, 0, s>>>(a);
    return cudaGetLastError();
}

cudaError_t launch_k3_choose(const K3Args &a, cudaStream_t s)
{
    const uint64_t tiles = k3_num_tiles(a.read_end - a.read_begin);
    if (tiles == 0) return cudaSuccess;
    const uint64_t chunks = a.pitch8 / 16;
    const unsigned g = (unsigned)tiles;
    if (chunks <= 1)       k3::choose_sparse_kernel<1><<<g, k3::THREADS, 0, s>>>(a);
    else if (chunks <= 2)  k3::choose_sparse_kernel<2><<<g, k3::THREADS, 0, s>>>(a);
    else if (chunks <= 4)  k3::choose_sparse_kernel<4><<<g, k3::THREADS, 0, s>>>(a);
    else if (chunks <= 8)  k3::choose_sparse_kernel<8><<<g, k3::THREADS, 0, s>>>(a);
    else if (chunks <= 32) k3::choose_wide_kernel<1><<<g, k3::THREADS, 0, s>>>(a);
    else if (chunks <= 64) k3::choose_wide_kernel<2><<<g, k3::THREADS, 0, s>>>(a);
    else                   k3::choose_kernel<32><<<g, k3::THREADS, 0, s>>>(a);   // more than 1024 genomes
    return cudaGetLastError();
}
cudaError_t launch_k3_export(const K3Args &a, cudaStream_t s)
{
    const uint64_t items = (uint64_t)(a.read_end - a.read_begin) * (a.pitch8 / 16);
    if (items == 0) return cudaSuccess;
    const uint64_t want = (items + (uint64_t)k3::THREADS * k3::EXPORT_UNROLL - 1) / ((uint64_t)k3::THREADS * k3::EXPORT_UNROLL);
    k3::export_kernel<<<(unsigned)std::min<uint64_t>(want, 148ull * 8), k3::THREADS, 0, s>>>(a);
    return cudaGetLastError();
}

// adds triples {read, genome << 8 | count} into uint8 cells (wrapping); keys unique within one call
cudaError_t launch_k3_import(const lime_pair_sim *tr, uint64_t n, uint8_t *sim8, uint64_t pitch8, uint32_t n_reads,
                             uint32_t n_refs, unsigned int *err, cudaStream_t s)
{
    if (n == 0) return cudaSuccess;
    const uint64_t want = (n + 255) / 256;
    k3::import_kernel<<<(unsigned)(want < 148ull * 32 ? want : 148ull * 32), 256, 0, s>>>(tr, n, sim8, pitch8, n_reads, n_refs, err);
    return cudaGetLastError();
}

cudaError_t launch_k3_emit(const K3Args &a, cudaStream_t s)
{
    const uint32_t n = a.read_end - a.read_begin;
    uint64_t want = ((uint64_t)n + k3::WARPS - 1) / k3::WARPS;
    if (want == 0) want = 1;                       // still writes the sentinel
    const unsigned grid = (unsigned)(want < 148ull * 16 ? want : 148ull * 16);
    k3::emit_kernel<<<grid, k3::THREADS, 0, s>>>(a);
    return cudaGetLastError();
}

} // namespace lime

// ---- clusterChoose without any allocation (lime_choose_static) ------------------------------------------------
// No counterpart in the reference: added for callers that cannot call
// cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it).  The scratch of the single-pass
// kernels -- the tile counter, the record count and one look-back word per tile of 256 reads -- lives in a
// file-scope device array, so every GPU that loads the library holds LIME_STATIC_SCRATCH_BYTES until exit.
// All calls share the array: it keeps the last call's contents (the part a call uses is cleared first, in stream
// order), and two calls that overlap on one GPU race.
namespace lime {
namespace k3 {
constexpr uint64_t STATIC_MAX_READS = LIME_CHOOSE_STATIC_MAX_READS;
constexpr uint64_t STATIC_MAX_TILES = (STATIC_MAX_READS + TILE_READS - 1) / TILE_READS;
constexpr uint64_t STATIC_HEAD = 256;                                   // bytes: [0] tile counter, [8] record count
constexpr uint64_t STATIC_BYTES = STATIC_HEAD + STATIC_MAX_TILES * 8;   // 512.25 KB
static __device__ __align__(256) unsigned char g_choose_scratch[STATIC_BYTES];
constexpr int STATIC_MAX_DEVICES = 64;
} // namespace k3
} // namespace lime

static void static_msg(char *err, uint64_t err_len, const char *what, const char *detail = nullptr)
{
    if (!err || !err_len) return;
    snprintf(err, (size_t)err_len, "lime_choose_static: %s%s%s", what, detail ? ": " : "", detail ? detail : "");
}

extern "C" int lime_choose_static(uint8_t *dev_sim, uint64_t pitch, uint32_t n_reads, uint32_t n_refs, uint32_t norm, float beta,
                                  lime_pair_sim *dev_bin, uint64_t bin_cap, uint64_t *dev_pos, uint64_t *dev_n_records,
                                  uint64_t *n_records, void *stream, char *err, uint64_t err_len)
{
    using namespace lime;
    if (err && err_len) err[0] = 0;
    if (n_reads > k3::STATIC_MAX_READS) {
        char d[128];
        snprintf(d, sizeof d, "%u reads exceed the static scratch (LIME_CHOOSE_STATIC_MAX_READS = %llu)", n_reads,
                 (unsigned long long)k3::STATIC_MAX_READS);
        static_msg(err, err_len, "size beyond the maximum", d);
        return LIME_ERR_ARG;
    }
    if (!dev_bin || bin_cap < 1 || (n_reads && (!dev_sim || !dev_pos)) || pitch % 16 || pitch < n_refs ||
        (reinterpret_cast<uintptr_t>(dev_sim) & 15) || (!dev_n_records && !n_records)) {
        static_msg(err, err_len, "bad argument (device pointers, pitch a multiple of 16 and >= n_refs, 16-byte aligned matrix, "
                                 "bin_cap >= 1, one of dev_n_records / n_records)");
        return LIME_ERR_ARG;
    }
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess || dev < 0 || dev >= k3::STATIC_MAX_DEVICES) {
        static_msg(err, err_len, "cudaGetDevice", e != cudaSuccess ? cudaGetErrorString(e) : "device ordinal out of range");
        return LIME_ERR_CUDA;
    }
    // the array's name is a host-side stand-in here: its device address, once per device, with that device current
    static unsigned char *scratch_of[k3::STATIC_MAX_DEVICES] = {nullptr};
    unsigned char *scratch = scratch_of[dev];
    if (!scratch) {
        e = cudaGetSymbolAddress(reinterpret_cast<void **>(&scratch), k3::g_choose_scratch);
        if (e != cudaSuccess || !scratch) {
            static_msg(err, err_len, "cudaGetSymbolAddress", cudaGetErrorString(e));
            return LIME_ERR_CUDA;
        }
        scratch_of[dev] = scratch;
    }
    const uint64_t tiles = k3_num_tiles(n_reads);
    // what must start at zero: the counters and the look-back words of the tiles this call uses
    e = cudaMemsetAsync(scratch, 0, k3::STATIC_HEAD + tiles * 8, s);
    if (e != cudaSuccess) { static_msg(err, err_len, "cudaMemsetAsync", cudaGetErrorString(e)); return LIME_ERR_CUDA; }
    unsigned long long *d_count = reinterpret_cast<unsigned long long *>(scratch + 8);
    if (tiles == 0) {
        // no reads: the sentinel record alone (ClusterBWT_DA.cpp:376-382)
        e = cudaMemsetAsync(dev_bin, 0, sizeof(lime_pair_sim), s);
        const unsigned long long one = 1;
        if (e == cudaSuccess) e = cudaMemcpyAsync(d_count, &one, 8, cudaMemcpyHostToDevice, s);
        if (e == cudaSuccess) e = cudaStreamSynchronize(s);          // `one` is a stack variable
    } else {
        K3Args a;
        a.sim8 = dev_sim; a.pitch8 = pitch; a.n_refs = n_refs;
        a.read_begin = 0; a.read_end = n_reads; a.norm = norm; a.beta = beta;
        a.tile_counter = reinterpret_cast<unsigned int *>(scratch);
        a.n_records = d_count;
        a.tile_state = reinterpret_cast<unsigned long long *>(scratch + k3::STATIC_HEAD);
        a.pos = dev_pos; a.bin = dev_bin; a.bin_cap = bin_cap;
        e = launch_k3_choose(a, s);
    }
    if (e != cudaSuccess) { static_msg(err, err_len, "launch", cudaGetErrorString(e)); return LIME_ERR_CUDA; }
    if (dev_n_records) {
        e = cudaMemcpyAsync(dev_n_records, d_count, 8, cudaMemcpyDeviceToDevice, s);
        if (e != cudaSuccess) { static_msg(err, err_len, "copy of the record count", cudaGetErrorString(e)); return LIME_ERR_CUDA; }
    }
    if (n_records) {
        unsigned long long total = 0;
        e = cudaMemcpyAsync(&total, d_count, 8, cudaMemcpyDeviceToHost, s);
        if (e == cudaSuccess) e = cudaStreamSynchronize(s);
        if (e != cudaSuccess) { static_msg(err, err_len, "read-back of the record count", cudaGetErrorString(e)); return LIME_ERR_CUDA; }
        *n_records = total;
        if (total > bin_cap) {
            static_msg(err, err_len, "dev_bin is too small for the records of this matrix (nothing was written past bin_cap)");
            return LIME_ERR_NOMEM;
        }
    }
    return LIME_OK;
}
