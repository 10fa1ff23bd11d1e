// lime_abi_load.cu -- lime_load(LIME_MEM_HOST) with a narrowed lcp transfer.
//
// The path reads lcp only through `lcp[i] < alpha` (ClusterLCP.cpp:150-202, :246-264).  When the caller
// has declared alpha <= 255 (lime_set_alpha_limit), the host arrays cross PCIe as 1 + 4 (+1) bytes per
// position instead of 4 + 4 (+1): host threads narrow lcp to min(lcp, 255) chunk by chunk into a ring
// of page-locked stage buffers (host_narrow.cpp) while the previous chunks are on the bus; the bytes stay on
// the device as they are: the kernels that read lcp (k12_stream.cu, template LCP8) take either width.
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
    ctx->in_lcp8.release();
}

namespace {

// the copy queue of narrow_transfer (narrow_loop.h) on the context's stream
struct CudaQueue {
    lime_ctx *ctx;
    const uint32_t *da; const uint8_t *ebwt;
    uint8_t *d8;
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
        cudaError_t e = cudaMemcpyAsync(ctx->in_da.as<uint32_t>() + base, da + base, len * 4, cudaMemcpyHostToDevice, ctx->stream);
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
    CudaQueue q{ctx, da, ebwt, ctx->in_lcp8.as<uint8_t>()};
    return narrow_transfer(q, lcp, n, chunk, threads, LIME_STAGE_SLOTS, ctx->stage);
}

extern "C" int lime_set_alpha_limit(lime_ctx *ctx, uint32_t max_alpha)
{
    if (!ctx) return LIME_ERR_ARG;
    ctx->alpha_limit = max_alpha;
    return LIME_OK;
}
