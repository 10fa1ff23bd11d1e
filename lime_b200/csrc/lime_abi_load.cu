// lime_abi_load.cu -- lime_load(LIME_MEM_HOST) with a narrowed lcp transfer.
//
// The path reads lcp only through `lcp[i] < alpha` (ClusterLCP.cpp:150-202, :246-264).  When the caller
// has declared alpha <= 255 (lime_set_alpha_limit), the host arrays cross PCIe as 1 + 4 (+1) bytes per
// position instead of 4 + 4 (+1): host threads narrow lcp to min(lcp, 255) chunk by chunk into a ring
// of page-locked stage buffers (host_narrow.cpp) while the previous chunks are on the bus; on the device
// widen_kernel writes the 32-bit array the step kernels read, so nothing downstream changes.
#include "abi_internal.cuh"
#include "narrow_loop.h"
#include <algorithm>

namespace lime {

// 16 narrowed positions per thread and iteration: one 16-byte load, four 16-byte stores
__global__ void __launch_bounds__(256) widen_kernel(const uint8_t *__restrict__ in, uint32_t *__restrict__ out, uint64_t n)
{
    const uint64_t n16 = n / 16;
    const uint4 *in16 = reinterpret_cast<const uint4 *>(in);
    uint4 *out4 = reinterpret_cast<uint4 *>(out);
    for (uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; i < n16; i += (uint64_t)gridDim.x * blockDim.x) {
        const uint4 v = __ldg(in16 + i);
        const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
        for (int k = 0; k < 4; k++)
            out4[4 * i + k] = make_uint4(w[k] & 255u, (w[k] >> 8) & 255u, (w[k] >> 16) & 255u, w[k] >> 24);
    }
    if (blockIdx.x == 0 && threadIdx.x < (n & 15)) out[n16 * 16 + threadIdx.x] = in[n16 * 16 + threadIdx.x];
}

cudaError_t launch_widen(const uint8_t *in, uint32_t *out, uint64_t n, cudaStream_t s)
{
    if (n == 0) return cudaSuccess;
    const uint64_t blocks = std::max<uint64_t>(1, std::min<uint64_t>((n / 16 + 255) / 256, 148 * 8));
    widen_kernel<<<(unsigned)blocks, 256, 0, s>>>(in, out, n);
    return cudaGetLastError();
}

} // namespace lime

extern "C" int lime_host_threads(void);

void load_release(lime_ctx *ctx)
{
    for (int s = 0; s < LIME_STAGE_SLOTS; s++) {
        if (ctx->stage_ev[s]) { cudaEventSynchronize(ctx->stage_ev[s]); cudaEventDestroy(ctx->stage_ev[s]); ctx->stage_ev[s] = nullptr; }
        if (ctx->stage[s]) { cudaFreeHost(ctx->stage[s]); ctx->stage[s] = nullptr; }
    }
    ctx->in_lcp8.release();
}

namespace {

// the copy queue of narrow_transfer (narrow_loop.h) on the context's stream
struct CudaQueue {
    lime_ctx *ctx;
    const uint32_t *da; const uint8_t *ebwt;
    uint8_t *d8; uint32_t *d32;
    int error(cudaError_t e, const char *what) { fail_cuda(ctx, e, what); return e == cudaErrorMemoryAllocation ? LIME_ERR_NOMEM : LIME_ERR_CUDA; }
    int stage_copy(int slot, uint64_t base, uint64_t len)
    {
        ctx->stage_busy = true;
        cudaError_t e = cudaMemcpyAsync(d8 + base, ctx->stage[slot], len, cudaMemcpyHostToDevice, ctx->stream);
        if (e != cudaSuccess) return error(e, "H2D lcp (narrowed)");
        e = cudaEventRecord(ctx->stage_ev[slot], ctx->stream);
        if (e != cudaSuccess) return error(e, "cudaEventRecord");
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
        cudaError_t e;
        {
            LaunchScope ls(ctx, KID_LOAD_WIDEN);
            e = launch_widen(d8 + base, d32 + base, len, ctx->stream);
        }
        if (e != cudaSuccess) return error(e, "load_widen launch");
        e = cudaMemcpyAsync(ctx->in_da.as<uint32_t>() + base, da + base, len * 4, cudaMemcpyHostToDevice, ctx->stream);
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
int load_host_narrow(lime_ctx *ctx, const uint32_t *lcp, const uint32_t *da, const uint8_t *ebwt, uint64_t n)
{
    const uint64_t chunk = LIME_STAGE_CHUNK;
    CK(ctx->in_lcp8.reserve((n + 15) / 16 * 16), "cudaMalloc narrowed lcp");
    for (int s = 0; s < LIME_STAGE_SLOTS; s++) {
        if (!ctx->stage[s]) CK(cudaHostAlloc((void **)&ctx->stage[s], chunk, cudaHostAllocDefault), "cudaHostAlloc lcp stage");
        if (!ctx->stage_ev[s]) CK(cudaEventCreateWithFlags(&ctx->stage_ev[s], cudaEventDisableTiming), "cudaEventCreate");
        else if (ctx->stage_busy) CK(cudaEventSynchronize(ctx->stage_ev[s]), "lcp stage of the previous load");
    }
    ctx->stage_busy = false;
    // the calling thread only queues copies (it yields while it waits): one core is left to it
    const int threads = std::max(1, lime_host_threads() - 1);
    CudaQueue q{ctx, da, ebwt, ctx->in_lcp8.as<uint8_t>(), ctx->in_lcp.as<uint32_t>()};
    return narrow_transfer(q, lcp, n, chunk, threads, LIME_STAGE_SLOTS, ctx->stage);
}

extern "C" int lime_set_alpha_limit(lime_ctx *ctx, uint32_t max_alpha)
{
    if (!ctx) return LIME_ERR_ARG;
    ctx->alpha_limit = max_alpha;
    return LIME_OK;
}
