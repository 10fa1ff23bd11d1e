// lime_abi_load.cu -- lime_load(LIME_MEM_HOST) with a narrowed lcp transfer.
//
// The path reads lcp only through `lcp[i] < alpha` (ClusterLCP.cpp:150-202, :246-264).  When the caller
// has declared alpha <= 255 (lime_set_alpha_limit), the host arrays cross PCIe as 1 + 4 (+1) bytes per
// position instead of 4 + 4 (+1): host threads narrow lcp to min(lcp, 255) chunk by chunk into a ring
// of page-locked stage buffers (host_narrow.cpp) while the previous chunks are on the bus; on the device
// widen_kernel writes the 32-bit array the step kernels read, so nothing downstream changes.
#include "abi_internal.cuh"
#include <algorithm>
#include <thread>

namespace lime {

struct NarrowPipe;
NarrowPipe *narrow_pipe_open(const uint32_t *src, uint64_t n, uint64_t chunk, int threads, int slots, uint8_t *const *stage);
bool narrow_pipe_ready(const NarrowPipe *p, uint64_t c);
void narrow_pipe_released(NarrowPipe *p, uint64_t k);
void narrow_pipe_close(NarrowPipe *p);

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
    const uint64_t n_chunks = (n + chunk - 1) / chunk;
    // the calling thread only queues copies; it yields while it waits
    const int threads = std::max(1, lime_host_threads() - 1);
    NarrowPipe *pipe = narrow_pipe_open(lcp, n, chunk, threads, LIME_STAGE_SLOTS, ctx->stage);
    struct Closer { NarrowPipe *p; ~Closer() { narrow_pipe_close(p); } } closer{pipe};
    uint8_t *d8 = ctx->in_lcp8.as<uint8_t>();
    uint32_t *d32 = ctx->in_lcp.as<uint32_t>();
    uint64_t queued = 0, released = 0;
    cudaError_t bad = cudaSuccess;
    auto poll = [&]() {
        while (released < queued) {
            const cudaError_t q = cudaEventQuery(ctx->stage_ev[released % LIME_STAGE_SLOTS]);
            if (q == cudaErrorNotReady) { cudaGetLastError(); break; }
            if (q != cudaSuccess) { bad = q; break; }
            released++;
        }
        narrow_pipe_released(pipe, released);
    };
    for (uint64_t c = 0; c < n_chunks; c++) {
        while (!narrow_pipe_ready(pipe, c)) {
            poll();
            if (bad != cudaSuccess) return fail_cuda(ctx, bad, "H2D lcp (narrowed)");
            std::this_thread::yield();
        }
        const uint64_t base = c * chunk, len = std::min(chunk, n - base);
        const int s = (int)(c % LIME_STAGE_SLOTS);
        ctx->stage_busy = true;
        CK(cudaMemcpyAsync(d8 + base, ctx->stage[s], len, cudaMemcpyHostToDevice, ctx->stream), "H2D lcp (narrowed)");
        CK(cudaEventRecord(ctx->stage_ev[s], ctx->stream), "cudaEventRecord");
        queued = c + 1;
        {
            LaunchScope ls(ctx, KID_LOAD_WIDEN);
            CK(launch_widen(d8 + base, d32 + base, len, ctx->stream), "load_widen launch");
        }
        CK(cudaMemcpyAsync(ctx->in_da.as<uint32_t>() + base, da + base, len * 4, cudaMemcpyHostToDevice, ctx->stream), "H2D da");
        if (ebwt) CK(cudaMemcpyAsync(ctx->in_ebwt.as<uint8_t>() + base, ebwt + base, len, cudaMemcpyHostToDevice, ctx->stream), "H2D ebwt");
        poll();
        if (bad != cudaSuccess) return fail_cuda(ctx, bad, "H2D lcp (narrowed)");
    }
    return LIME_OK;
}

extern "C" int lime_set_alpha_limit(lime_ctx *ctx, uint32_t max_alpha)
{
    if (!ctx) return LIME_ERR_ARG;
    ctx->alpha_limit = max_alpha;
    return LIME_OK;
}
