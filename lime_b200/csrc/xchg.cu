// xchg.cu -- multi-GPU routing of similarity updates (the exchange step of north_star: partial (read, genome,
// count) triples go to the GPU that owns the read).
//
// With the position range sharded over GPUs, a GPU meets reads of every owner.  The fused kernel
// (k12_stream.cu, MODE_FLAT) and the block path append every update `SimArray_[r][g] += t`
// (ClusterBWT_DA.cpp:183,247) as a record {r, g << 8 | t} to a flat list in local HBM.  Then
//   route_kernel  streams that list once, sorts each tile of 2048 records by owner in shared memory and
//                 writes every owner's run with coalesced stores straight into that owner's inbox -- peer
//                 memory over NVLink / NVSwitch (CUDA IPC mapping), or local memory for its own reads.
//                 Space in an inbox region (one per source GPU) is taken with a LOCAL atomic: only this GPU
//                 writes to its region, so nothing remote is ever read or updated atomically.
//                 Inside an owner the records are also split by row block (n_sub blocks of the slab, each small
//                 enough for L2), so that the owner can apply block after block with its slab lines in L2.
//   publish_kernel (behind route_kernel in stream order) stores this source's record counts into a mailbox at the
//                 head of every owner's inbox and then raises a flag there (the number of the accumulation);
//   wait_kernel   (on the owner) one warp polls the flags of all sources -- bounded, 20 s -- so that the kernels
//                 queued behind it see every region complete;
//   apply_kernel  adds the records of one row block from every source region to the owner's uint8 rows with
//                 fire-and-forget adds and sums their t for the carry proof; one launch per row block.
// No collective and no host synchronisation is involved: records, counts and flags all travel as stores to peer
// memory, so the whole chain can be queued on a side stream and runs beside the scoring of the next strand.
#include "common.cuh"
#include <algorithm>
#include <cstdlib>

namespace lime {
namespace xchg {

constexpr int THREADS = 256;                    // apply_kernel
constexpr int PER_THREAD = 8;                   // route_kernel<T>: records per thread and tile (tile = T * 8 records)

constexpr int BINS_MAX = XCHG_MAX * XCHG_SUB_MAX;
struct RouteArgs {
    const uint2 *flat;
    const unsigned long long *flat_count;       // slots in use (device); clamped to flat_cap
    uint64_t flat_cap;
    uint32_t world, per, magic;                 // owner of read r = r / per; magic = min(floor(2^32 / per), 2^32 - 1)
    uint32_t n_sub, sub_rows, sub_magic;        // row block of a row inside its owner's slab = row / sub_rows
    uint32_t n_refs;
    uint2 *dst[XCHG_MAX];                       // owner o's inbox region for this source and strand: [n_sub][sub_cap]
    uint64_t sub_cap;                           // records per row block of a region
    unsigned long long *sent;                   // [world][n_sub] records written (device, local)
    unsigned long long *err;                    // bit 2: a region overflowed, bit 3: record out of range
};

__device__ __forceinline__ uint32_t owner_of(uint32_t r, uint32_t per, uint32_t magic, uint32_t &rem)
{
    uint32_t q = __umulhi(r, magic);            // quotient or quotient - 1
    rem = r - q * per;
    if (rem >= per) { q++; rem -= per; }
    return q;
}

template <int T>
__global__ void __launch_bounds__(T) route_kernel(const RouteArgs P)
{
    constexpr int TILE = T * PER_THREAD;
    constexpr int THREADS = T;
    extern __shared__ uint2 s_rec[];             // TILE records (dynamic: 64 KB at 1024 threads)
    __shared__ unsigned s_cnt[BINS_MAX], s_off[BINS_MAX + 1];
    __shared__ unsigned long long s_base[BINS_MAX];
    const unsigned tid = threadIdx.x, lane = tid & 31u;
    const unsigned n_bins = P.world * P.n_sub;
    const unsigned long long n = min(*P.flat_count, (unsigned long long)P.flat_cap);
    const unsigned long long n_tiles = (n + TILE - 1) / TILE;
    for (unsigned long long tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
        if (tid < BINS_MAX) s_cnt[tid] = 0;
        __syncthreads();
        uint2 rec[PER_THREAD];
        unsigned own[PER_THREAD], idx[PER_THREAD];
#pragma unroll
        for (int k = 0; k < PER_THREAD; k++) {
            const unsigned long long i = tile * TILE + (unsigned long long)k * THREADS + tid;
            rec[k] = i < n ? __ldcs(P.flat + i) : make_uint2(0u, 0u);
        }
#pragma unroll
        for (int k = 0; k < PER_THREAD; k++) {
            bool valid = (rec[k].y & 0xffu) != 0u;                     // t = 0: padding of a warp's last block
            uint32_t rem = 0, q = 0;
            if (valid) {
                q = owner_of(rec[k].x, P.per, P.magic, rem);
                if (q >= P.world || (rec[k].y >> 8) >= P.n_refs) { valid = false; atomicOr(P.err, 8ull); }
                else {
                    uint32_t r2;
                    q = q * P.n_sub + owner_of(rem, P.sub_rows, P.sub_magic, r2);      // bin = (owner, row block)
                }
            }
            rec[k].x = rem;                                            // row inside the owner's slab
            // the lanes that go to the same bin take their slots with one shared-memory atomic
            const unsigned mm = __match_any_sync(FULL, valid ? q : 0xffffffffu);
            unsigned base = 0;
            const int leader = __ffs(mm) - 1;
            if (valid && (int)lane == leader) base = atomicAdd(&s_cnt[q], (unsigned)__popc(mm));
            base = __shfl_sync(FULL, base, leader);
            own[k] = valid ? q : 0xffffffffu;
            idx[k] = base + __popc(mm & ((1u << lane) - 1u));
        }
        __syncthreads();
        if (tid < 32) {
            // exclusive prefix of the bin counts (n_bins <= 128: four per lane)
            unsigned c4[4], sum = 0;
#pragma unroll
            for (int j = 0; j < 4; j++) { const unsigned b = 4 * tid + j; c4[j] = b < n_bins ? s_cnt[b] : 0u; sum += c4[j]; }
            unsigned incl = sum;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const unsigned y = __shfl_up_sync(FULL, incl, o);
                if (tid >= (unsigned)o) incl += y;
            }
            unsigned run = incl - sum;
#pragma unroll
            for (int j = 0; j < 4; j++) { const unsigned b = 4 * tid + j; if (b < n_bins) s_off[b] = run; run += c4[j]; }
            if (tid == 31) s_off[n_bins] = incl;
        }
        if (tid >= 64 && tid - 64 < n_bins && s_cnt[tid - 64])
            s_base[tid - 64] = atomicAdd(&P.sent[tid - 64], (unsigned long long)s_cnt[tid - 64]);
        __syncthreads();
#pragma unroll
        for (int k = 0; k < PER_THREAD; k++)
            if (own[k] != 0xffffffffu) s_rec[s_off[own[k]] + idx[k]] = rec[k];
        __syncthreads();
        // sorted by bin: a warp takes a bin's run and writes it with 16-byte stores (two records each; a single
        // record in front when the destination starts on an odd slot, one behind when the rest is odd)
        for (unsigned b = tid >> 5; b < n_bins; b += THREADS / 32) {
            const unsigned first = s_off[b], cnt = s_off[b + 1] - first;
            if (!cnt) continue;
            const unsigned long long d0 = s_base[b];
            unsigned m = cnt;
            if (d0 + cnt > P.sub_cap) {                               // region full: keep what fits, report
                m = d0 < P.sub_cap ? (unsigned)(P.sub_cap - d0) : 0u;
                if (lane == 0) atomicOr(P.err, 4ull);
            }
            const unsigned o = b / P.n_sub, sb = b - o * P.n_sub;
            uint2 *dst = P.dst[o] + (size_t)sb * P.sub_cap + d0;
            const uint2 *src = s_rec + first;
            const unsigned head = (unsigned)(d0 & 1ull) < m ? (unsigned)(d0 & 1ull) : m;
            if (lane == 0 && head) dst[0] = src[0];
            const unsigned pairs = (m - head) >> 1;
            for (unsigned k = lane; k < pairs; k += 32) {
                const uint2 a = src[head + 2 * k], c = src[head + 2 * k + 1];
                *reinterpret_cast<uint4 *>(dst + head + 2 * k) = make_uint4(a.x, a.y, c.x, c.y);
            }
            if (lane == 0 && ((m - head) & 1u)) dst[m - 1] = src[m - 1];
        }
        __syncthreads();
    }
    __threadfence_system();                     // the stores to peer memory are performed before the kernel ends
}

__global__ void publish_kernel(const XchgPublish P)
{
    const unsigned o = threadIdx.x;
    if (o >= P.world) return;
    volatile unsigned long long *m = P.mail[o];
    for (unsigned j = 0; j < P.n_sub; j++) m[j] = P.sent[o * P.n_sub + j];
    // route_kernel has completed (stream order) and fenced its stores; the counts before the flag
    __threadfence_system();
    m[XCHG_SUB_MAX] = P.epoch;
    __threadfence_system();
}

__device__ __forceinline__ unsigned long long global_ns()
{
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}

__global__ void wait_kernel(const XchgWait P)
{
    const unsigned p = threadIdx.x;
    if (p >= P.world || (P.only >= 0 && (int)p != P.only)) return;
    const volatile unsigned long long *flag = P.mail + (size_t)p * XCHG_MAIL_STRIDE + XCHG_SUB_MAX;
    const unsigned long long t0 = global_ns();
    while (*flag < P.epoch) {
        __nanosleep(200);
        if (global_ns() - t0 > 20000000000ull) { atomicOr(P.err, 16ull); break; }   // a source never delivered
    }
    __threadfence_system();
}

struct ApplyArgs {
    const uint2 *src;                           // this owner's inbox regions of one strand: [world][n_sub][sub_cap]
    uint64_t sub_cap;
    const unsigned long long *mail;             // [world][XCHG_MAIL_STRIDE]: records sent by source p for row block j
    int only;
    uint32_t world, rank, n_sub, sub;           // sub: the row block this launch applies
    uint8_t *slab; uint64_t pitch;
    uint32_t rows, n_refs;
    unsigned long long *applied_t;              // sum of t added (carry proof: must equal the byte sum of the slab)
    unsigned long long *applied_n;
    unsigned long long *err;                    // bit 2: a source sent more than a region holds, bit 3: bad record
};

// gridDim.y = sources
__global__ void __launch_bounds__(THREADS) apply_kernel(const ApplyArgs P)
{
    const unsigned p = blockIdx.y;
    if (P.only >= 0 && (int)p != P.only) return;
    unsigned long long n = P.mail[(size_t)p * XCHG_MAIL_STRIDE + P.sub];
    if (n > P.sub_cap) { if (threadIdx.x == 0 && blockIdx.x == 0) atomicOr(P.err, 4ull); n = P.sub_cap; }
    const uint2 *src = P.src + ((size_t)p * P.n_sub + P.sub) * P.sub_cap;
    uint64_t pol;
    asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(pol));
    unsigned long long tsum = 0, cnt = 0;
    // two records per 16-byte load; regions are 16-byte aligned
    const unsigned long long n2 = n / 2;
    const uint4 *src2 = reinterpret_cast<const uint4 *>(src);
    for (unsigned long long i = (unsigned long long)blockIdx.x * THREADS + threadIdx.x; i < n2 + (n & 1ull);
         i += (unsigned long long)gridDim.x * THREADS) {
        uint2 r[2];
        int m = 2;
        if (i < n2) { const uint4 v = *(src2 + i); r[0] = make_uint2(v.x, v.y); r[1] = make_uint2(v.z, v.w); }
        else { r[0] = src[n - 1]; r[1] = make_uint2(0u, 0u); m = 1; }
#pragma unroll
        for (int k = 0; k < 2; k++) {
            if (k >= m) continue;
            const uint32_t row = r[k].x, g = r[k].y >> 8, t = r[k].y & 0xffu;
            if (row >= P.rows || g >= P.n_refs) { atomicOr(P.err, 8ull); continue; }
            unsigned *w = reinterpret_cast<unsigned *>(P.slab + (size_t)row * P.pitch) + (g >> 2);
            asm volatile("red.relaxed.gpu.global.add.L2::cache_hint.u32 [%0], %1, %2;" ::"l"(w), "r"(t << ((g & 3u) * 8u)), "l"(pol) : "memory");
            tsum += t; cnt++;
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) { tsum += __shfl_xor_sync(FULL, tsum, o); cnt += __shfl_xor_sync(FULL, cnt, o); }
    __shared__ unsigned long long s_t[THREADS / 32], s_n[THREADS / 32];
    if ((threadIdx.x & 31) == 0) { s_t[threadIdx.x >> 5] = tsum; s_n[threadIdx.x >> 5] = cnt; }
    __syncthreads();
    if (threadIdx.x == 0) {
        unsigned long long a = 0, b = 0;
        for (int k = 0; k < THREADS / 32; k++) { a += s_t[k]; b += s_n[k]; }
        if (a) atomicAdd(P.applied_t, a);
        if (b) atomicAdd(P.applied_n, b);
    }
}

} // namespace xchg

cudaError_t launch_xchg_route(const XchgRoute &a, cudaStream_t s)
{
    xchg::RouteArgs P;
    P.flat = a.flat; P.flat_count = a.flat_count; P.flat_cap = a.flat_cap;
    P.world = a.world; P.per = a.per;
    P.magic = (uint32_t)std::min<uint64_t>((1ull << 32) / std::max<uint32_t>(a.per, 1u), 0xffffffffull);
    P.n_sub = a.n_sub; P.sub_rows = a.sub_rows;
    P.sub_magic = (uint32_t)std::min<uint64_t>((1ull << 32) / std::max<uint32_t>(a.sub_rows, 1u), 0xffffffffull);
    P.n_refs = a.n_refs;
    for (int o = 0; o < XCHG_MAX; o++) P.dst[o] = o < (int)a.world ? a.dst[o] : nullptr;
    P.sub_cap = a.sub_cap; P.sent = a.sent; P.err = a.err;
    // tile = 8 records per thread; the longer the tile, the longer the run a bin gets (stores over NVLink want
    // hundreds of bytes): 512 threads -> 4096 records, 32 KB of shared memory, 2-3 CTAs per SM
    static int threads = 0;
    if (!threads) {
        threads = 512;
        if (const char *e = getenv("LIME_XCHG_ROUTE_THREADS")) threads = atoi(e);
    }
    if (threads == 256) xchg::route_kernel<256><<<148 * 4, 256, 256 * 64, s>>>(P);
    else if (threads == 1024) {
        static bool once = false;
        if (!once) {
            const cudaError_t e = cudaFuncSetAttribute(xchg::route_kernel<1024>, cudaFuncAttributeMaxDynamicSharedMemorySize, 1024 * 64);
            if (e != cudaSuccess) return e;
            once = true;
        }
        xchg::route_kernel<1024><<<148 * 2, 1024, 1024 * 64, s>>>(P);
    } else xchg::route_kernel<512><<<148 * 3, 512, 512 * 64, s>>>(P);
    return cudaGetLastError();
}

cudaError_t launch_xchg_apply(const XchgApply &a, cudaStream_t s)
{
    xchg::ApplyArgs P;
    P.src = a.src; P.sub_cap = a.sub_cap; P.mail = a.mail; P.only = a.only; P.world = a.world; P.rank = a.rank;
    P.n_sub = a.n_sub; P.sub = a.sub;
    P.slab = a.slab; P.pitch = a.pitch; P.rows = a.rows; P.n_refs = a.n_refs;
    P.applied_t = a.applied_t; P.applied_n = a.applied_n; P.err = a.err;
    const dim3 grid((148 * 8 + a.world - 1) / a.world, a.world);
    xchg::apply_kernel<<<grid, xchg::THREADS, 0, s>>>(P);
    return cudaGetLastError();
}

cudaError_t launch_xchg_publish(const XchgPublish &a, cudaStream_t s)
{
    xchg::publish_kernel<<<1, 32, 0, s>>>(a);
    return cudaGetLastError();
}

cudaError_t launch_xchg_wait(const XchgWait &a, cudaStream_t s)
{
    xchg::wait_kernel<<<1, 32, 0, s>>>(a);
    return cudaGetLastError();
}

} // namespace lime
