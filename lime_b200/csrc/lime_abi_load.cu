// lime_abi_load.cu -- lime_load(LIME_MEM_HOST) with a narrowed lcp transfer.
//
// The path reads lcp only through `lcp[i] < alpha` (ClusterLCP.cpp:150-202, :246-264).  When the caller
// has declared alpha <= 255 (lime_set_alpha_limit), the host arrays cross PCIe as 1 + 4 (+1) bytes per
// position instead of 4 + 4 (+1): host threads narrow lcp to min(lcp, 255) chunk by chunk into a ring
// of page-locked stage buffers (host_narrow.cpp) while the previous chunks are on the bus; the bytes stay on
// the device as they are: the kernels that read lcp (k12_stream.cu, template LCP8) take either width.
// When the caller has named the alpha itself (lime_set_alpha), one BIT per position crosses instead -- lcp < alpha,
// packed by the same host threads -- and expand_kernel turns every chunk into the bytes 0 (head) / 255 behind its
// copy; the kernels see an LCP8 array that answers lcp < alpha for that alpha only (the steps check it).  In that
// mode da crosses as 3 bytes per position while every document id fits (n_reads + n_refs <= 2^24; packed by the same
// threads, expanded by expand3_kernel into the uint32 array the kernels read): 1/8 + 3 (+1) bytes per position on
// the bus instead of 4 + 4 (+1).
#include "abi_internal.cuh"
#include "narrow_loop.h"
#include <algorithm>

extern "C" int lime_host_threads(void);

void load_release(lime_ctx *ctx)
{
    for (int s = 0; s < LIME_STAGE_SLOTS; s++) {
        if (ctx->stage_ev[s]) { cudaEventSynchronize(ctx->stage_ev[s]); cudaEventDestroy(ctx->stage_ev[s]); ctx->stage_ev[s] = nullptr; }
        if (ctx->stage[s]) { cudaFreeHost(ctx->stage[s]); ctx->stage[s] = nullptr; }
    }
    ctx->in_lcp8.release(); ctx->in_lcpbits.release(); ctx->in_da3.release();
    ctx->stage_bytes = 0;
}

namespace lime {
// out[8 i + j] = bit j of bits[i] ? 0 : 255 for the n positions of a chunk: a thread expands two bytes into 16
__global__ void __launch_bounds__(256) expand_kernel(const uint8_t *__restrict__ bits, uint8_t *__restrict__ out, uint64_t n)
{
    const uint64_t t = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const uint64_t p = t * 16;
    if (p >= n) return;
    const unsigned b = (unsigned)bits[2 * t] | (p + 8 < n ? (unsigned)bits[2 * t + 1] << 8 : 0xff00u);
    uint32_t w[4];
#pragma unroll
    for (int k = 0; k < 4; k++) {
        const unsigned nib = (b >> (4 * k)) & 15u;
        // bit set -> byte 0x00, clear -> 0xff: spread the 4 bits to the low bit of each byte, then 0xff * (1 - bit)
        const uint32_t spread = (nib & 1u) | ((nib & 2u) << 7) | ((nib & 4u) << 14) | ((nib & 8u) << 21);
        w[k] = (0x01010101u - spread) * 0xffu;
    }
    if (p + 16 <= n) *reinterpret_cast<uint4 *>(out + p) = make_uint4(w[0], w[1], w[2], w[3]);
    else
        for (uint64_t e = 0; p + e < n; e++) out[p + e] = (uint8_t)(w[e >> 2] >> (8 * (e & 3)));
}
// out[i] = the three bytes in[3 i .. 3 i + 2], little endian (document ids below 2^24): a thread takes 4 positions =
// three 32-bit words in, one 16-byte store out; `in` is 4-byte aligned and readable up to 3 n rounded up to 12
__global__ void __launch_bounds__(256) expand3_kernel(const uint32_t *__restrict__ in, uint32_t *__restrict__ out, uint64_t n)
{
    const uint64_t t = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const uint64_t p = t * 4;
    if (p >= n) return;
    const uint32_t a = in[3 * t], b = in[3 * t + 1], c = in[3 * t + 2];
    const uint32_t v[4] = {a & 0xffffffu, (a >> 24) | ((b & 0xffffu) << 8), (b >> 16) | ((c & 0xffu) << 16), c >> 8};
    if (p + 4 <= n) *reinterpret_cast<uint4 *>(out + p) = make_uint4(v[0], v[1], v[2], v[3]);
    else
        for (uint64_t e = 0; p + e < n; e++) out[p + e] = v[e];
}
} // namespace lime

namespace {

// the copy queue of narrow_transfer (narrow_loop.h) on the context's stream
struct CudaQueue {
    lime_ctx *ctx;
    const uint32_t *da; const uint8_t *ebwt;
    uint8_t *d8;
    uint8_t *dbits;       // non-null: the stage holds one bit per position; expanded into d8 behind the copy
    uint8_t *dda3;        // non-null: the stage also holds da as 3 bytes per position (at stage + chunk); expanded into in_da
    uint64_t chunk;
    int error(cudaError_t e, const char *what) { fail_cuda(ctx, e, what); return e == cudaErrorMemoryAllocation ? LIME_ERR_NOMEM : LIME_ERR_CUDA; }
    int stage_copy(int slot, uint64_t base, uint64_t len)
    {
        ctx->stage_busy = true;
        cudaError_t e;
        if (dbits) e = cudaMemcpyAsync(dbits + base / 8, ctx->stage[slot], (len + 7) / 8, cudaMemcpyHostToDevice, ctx->stream);
        else e = cudaMemcpyAsync(d8 + base, ctx->stage[slot], len, cudaMemcpyHostToDevice, ctx->stream);
        if (e != cudaSuccess) return error(e, "H2D lcp (narrowed)");
        if (dda3) {
            e = cudaMemcpyAsync(dda3 + 3 * base, ctx->stage[slot] + chunk, 3 * len, cudaMemcpyHostToDevice, ctx->stream);
            if (e != cudaSuccess) return error(e, "H2D da (3 bytes per position)");
        }
        e = cudaEventRecord(ctx->stage_ev[slot], ctx->stream);
        if (e != cudaSuccess) return error(e, "cudaEventRecord");
        if (dbits) {
            const uint64_t threads = (len + 15) / 16;
            ctx->launches++;
            lime::expand_kernel<<<(unsigned)((threads + 255) / 256), 256, 0, ctx->stream>>>(dbits + base / 8, d8 + base, len);
            e = cudaGetLastError();
            if (e != cudaSuccess) return error(e, "expand launch");
        }
        if (dda3) {
            const uint64_t threads = (len + 3) / 4;
            ctx->launches++;
            lime::expand3_kernel<<<(unsigned)((threads + 255) / 256), 256, 0, ctx->stream>>>(
                reinterpret_cast<const uint32_t *>(dda3 + 3 * base), ctx->in_da.as<uint32_t>() + base, len);
            e = cudaGetLastError();
            if (e != cudaSuccess) return error(e, "expand3 launch");
        }
        return 0;
    }
    int slot_state(int slot)
    {
        const cudaError_t e = cudaEventQuery(ctx->stage_ev[slot]);
        if (e == cudaSuccess) return 0;
        if (e == cudaErrorNotReady) { cudaGetLastError(); return 1; }
        return error(e, "H2D lcp (narrowed)");
    }
    int rest(uint64_t base, uint64_t len)
    {
        cudaError_t e = cudaSuccess;
        if (!dda3) e = cudaMemcpyAsync(ctx->in_da.as<uint32_t>() + base, da + base, len * 4, cudaMemcpyHostToDevice, ctx->stream);
        if (e != cudaSuccess) return error(e, "H2D da");
        if (ebwt) {
            e = cudaMemcpyAsync(ctx->in_ebwt.as<uint8_t>() + base, ebwt + base, len, cudaMemcpyHostToDevice, ctx->stream);
            if (e != cudaSuccess) return error(e, "H2D ebwt");
        }
        return 0;
    }
};

} // namespace

// lcp, da (and ebwt) of n positions: host -> device, lcp narrowed.  The device arrays are reserved by the caller.
int load_host_narrow(lime_ctx *ctx, const uint32_t *lcp, const uint32_t *da, const uint8_t *ebwt, uint64_t n, uint32_t bits_alpha,
                     bool da3)
{
    const uint64_t chunk = LIME_STAGE_CHUNK;
    // a stage buffer holds the lcp bytes / bits of a chunk and, with da3, its da as 3 bytes per position behind them
    const uint64_t stage_bytes = da3 ? 4 * chunk : chunk;
    if (ctx->stage_bytes < stage_bytes) {
        for (int s = 0; s < LIME_STAGE_SLOTS; s++) {
            if (ctx->stage_ev[s] && ctx->stage_busy) CK(cudaEventSynchronize(ctx->stage_ev[s]), "lcp stage of the previous load");
            if (ctx->stage[s]) { cudaFreeHost(ctx->stage[s]); ctx->stage[s] = nullptr; }
        }
        ctx->stage_busy = false;
        ctx->stage_bytes = stage_bytes;
    }
    if (da3) CK(ctx->in_da3.reserve(3 * n + 32), "cudaMalloc da (3 bytes per position)");
    CK(ctx->in_lcp8.reserve((n + 15) / 16 * 16), "cudaMalloc narrowed lcp");
    if (bits_alpha) CK(ctx->in_lcpbits.reserve((n + 7) / 8 + 16), "cudaMalloc lcp bits");
    for (int s = 0; s < LIME_STAGE_SLOTS; s++) {
        if (!ctx->stage[s]) CK(cudaHostAlloc((void **)&ctx->stage[s], ctx->stage_bytes, cudaHostAllocDefault), "cudaHostAlloc lcp stage");
        if (!ctx->stage_ev[s]) CK(cudaEventCreateWithFlags(&ctx->stage_ev[s], cudaEventDisableTiming), "cudaEventCreate");
        else if (ctx->stage_busy) CK(cudaEventSynchronize(ctx->stage_ev[s]), "lcp stage of the previous load");
    }
    ctx->stage_busy = false;
    // the calling thread only queues copies (it yields while it waits): one core is left to it
    const int threads = std::max(1, lime_host_threads() - 1);
    CudaQueue q{ctx, da, ebwt, ctx->in_lcp8.as<uint8_t>(), bits_alpha ? ctx->in_lcpbits.as<uint8_t>() : nullptr,
                da3 ? ctx->in_da3.as<uint8_t>() : nullptr, chunk};
    return narrow_transfer(q, lcp, n, chunk, threads, LIME_STAGE_SLOTS, ctx->stage, bits_alpha, da3 ? da : nullptr);
}

extern "C" int lime_set_alpha_limit(lime_ctx *ctx, uint32_t max_alpha)
{
    if (!ctx) return LIME_ERR_ARG;
    ctx->alpha_limit = max_alpha;
    return LIME_OK;
}

extern "C" int lime_set_alpha(lime_ctx *ctx, uint32_t alpha)
{
    if (!ctx) return LIME_ERR_ARG;
    ctx->alpha_exact = alpha;
    return LIME_OK;
}
