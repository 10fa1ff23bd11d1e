// lime_abi_xchg.cu -- multi-GPU accumulation: similarity rows live on the GPU that owns the read (one process
// per GPU, one NVLink / NVSwitch node), updates travel there as records (kernels: k12_stream.cu MODE_FLAT,
// xchg.cu).
//
// The position range of a strand is sharded over the GPUs, the reads are owned by contiguous ranges.  Every
// GPU keeps only the rows of ITS reads (a slab per strand) and an inbox with one region per source GPU, which
// the other processes map through CUDA IPC.  Per strand and GPU:
//   lime_xchg_accumulate   fused kernel + block path append {read, genome << 8 | t} records to a local list;
//                          route_kernel sorts them by owner and writes each owner's run into this GPU's region
//                          of that owner's inbox (bulk stores over NVLink); the per-owner record counts are left
//                          in device memory for the caller's all-gather -- the job's only collective per strand,
//                          which is also the barrier "every source has delivered";
//   lime_xchg_apply        adds the records of all regions to the slab (local fire-and-forget adds);
//   lime_xchg_choose       clusterChoose on the slab (also: byte sum for the carry proof, rows left zero).
// None of these synchronises with the host; lime_classify reads the strands' state back and reports
// LIME_ERR_RETRY when a strand must be redone through the exact route (a cell above 255, a list too small).
#include "abi_internal.cuh"
#include <algorithm>
#include <cstring>

static_assert(sizeof(cudaIpcMemHandle_t) == LIME_IPC_HANDLE_BYTES, "IPC handle size");
static_assert(XCHG_MAX == LIME_XCHG_MAX, "exchange limit");

void xchg_release(lime_ctx *ctx)
{
    XchgState &S = ctx->xchg;
    for (int k = 0; k < XCHG_MAX; k++) {
        if (S.inbox_map[k] && S.inbox_map[k] != S.inbox_own) cudaIpcCloseMemHandle(S.inbox_map[k]);
        S.inbox_map[k] = nullptr;
    }
    if (S.inbox_own) cudaFree(S.inbox_own);
    S.inbox_own = nullptr;
    S.slabs.release(); S.flat.release();
    if (S.active) { ctx->d_sim = nullptr; ctx->sim_rows = 0; ctx->sim_clean = false; }
    S.active = false; S.world = 0; S.n_strands = 0; S.loopback = false;
}

extern "C" {

int lime_xchg_alloc(lime_ctx *ctx, int world, int rank, uint32_t reads_per_rank, uint32_t n_refs, int n_strands,
                    uint64_t flat_cap, uint64_t inbox_cap, uint8_t *ipc_handle)
{
    if (!ctx || !ipc_handle || world < 1 || world > LIME_XCHG_MAX || rank < 0 || rank >= world ||
        n_strands < 1 || n_strands > LIME_MAXFILES || reads_per_rank == 0)
        return LIME_ERR_ARG;
    if (n_refs >= (1u << 24)) return fail(ctx, LIME_ERR_ARG, "lime_xchg_alloc: a record holds genome << 8 | t, n_refs must be below 2^24");
    CK(cudaSetDevice(ctx->device), "cudaSetDevice");
    CK(cudaStreamSynchronize(ctx->stream), "lime_xchg_alloc");
    xchg_release(ctx);
    XchgState &S = ctx->xchg;
    S.world = world; S.rank = rank; S.n_strands = n_strands; S.per = reads_per_rank;
    S.rows = reads_per_rank; S.pitch = lime_sim_pitch(n_refs);
    S.flat_cap = std::max<uint64_t>((flat_cap + K12_FLAT_BLK - 1) / K12_FLAT_BLK * K12_FLAT_BLK, K12_FLAT_BLK);
    // the slab is applied in row blocks of at most 32 MB, so that its lines stay in L2 while their records arrive
    S.n_sub = (uint32_t)std::min<uint64_t>(XCHG_SUB_MAX, std::max<uint64_t>(1, (S.rows * S.pitch + (32ull << 20) - 1) / (32ull << 20)));
    S.sub_rows = (uint32_t)((S.rows + S.n_sub - 1) / S.n_sub);
    S.sub_cap = std::max<uint64_t>(((inbox_cap + S.n_sub - 1) / S.n_sub * 5 / 4 + 255) / 256 * 256, 256);
    S.inbox_cap = S.sub_cap * S.n_sub;
    const uint64_t slab_bytes = std::max<uint64_t>(S.rows * S.pitch * (uint64_t)n_strands, 16);
    CK(S.slabs.reserve(slab_bytes), "cudaMalloc slabs");
    CK(cudaMemsetAsync(S.slabs.p, 0, slab_bytes, ctx->stream), "memset slabs");
    CK(S.flat.reserve(S.flat_cap * sizeof(uint2)), "cudaMalloc update list");
    const uint64_t inbox_bytes = S.inbox_cap * sizeof(uint2) * (uint64_t)world * (uint64_t)n_strands;
    CK(cudaMalloc(&S.inbox_own, inbox_bytes), "cudaMalloc inbox");          // plain cudaMalloc: IPC needs it
    CK(ctx->strand_state.reserve(sizeof(unsigned long long) * SS_COUNT * (LIME_MAXFILES + 1)), "cudaMalloc strand state");
    if (!ctx->strand_host)
        CK(cudaHostAlloc((void **)&ctx->strand_host, sizeof(unsigned long long) * SS_COUNT * (LIME_MAXFILES + 1), cudaHostAllocDefault),
           "cudaHostAlloc strand state");
    cudaIpcMemHandle_t h;
    CK(cudaIpcGetMemHandle(&h, S.inbox_own), "cudaIpcGetMemHandle");
    memcpy(ipc_handle, &h, sizeof(h));
    CK(cudaStreamSynchronize(ctx->stream), "lime_xchg_alloc");
    return LIME_OK;
}

int lime_xchg_open(lime_ctx *ctx, const uint8_t *ipc_handles)
{
    if (!ctx) return LIME_ERR_ARG;
    XchgState &S = ctx->xchg;
    if (!S.inbox_own) return fail(ctx, LIME_ERR_STATE, "lime_xchg_open: lime_xchg_alloc first");
    CK(cudaSetDevice(ctx->device), "cudaSetDevice");
    S.loopback = ipc_handles == nullptr;
    for (int k = 0; k < S.world; k++) {
        if (k == S.rank || S.loopback) { S.inbox_map[k] = S.inbox_own; continue; }
        cudaIpcMemHandle_t h;
        memcpy(&h, ipc_handles + (size_t)k * LIME_IPC_HANDLE_BYTES, sizeof(h));
        CK(cudaIpcOpenMemHandle(&S.inbox_map[k], h, cudaIpcMemLazyEnablePeerAccess), "cudaIpcOpenMemHandle");
    }
    S.active = true;
    ctx->sim_bound = false; ctx->d_sim = nullptr; ctx->sim_rows = 0; ctx->sim_pitch = S.pitch; ctx->sim_clean = false;
    return LIME_OK;
}

int lime_xchg_counts_len(lime_ctx *ctx)
{
    if (!ctx || !ctx->xchg.inbox_own) return 0;
    return ctx->xchg.world * (int)ctx->xchg.n_sub;
}

int lime_xchg_close(lime_ctx *ctx)
{
    if (!ctx) return LIME_ERR_ARG;
    CK(cudaSetDevice(ctx->device), "cudaSetDevice");
    CK(cudaStreamSynchronize(ctx->stream), "lime_xchg_close");
    xchg_release(ctx);
    return LIME_OK;
}

int lime_xchg_accumulate(lime_ctx *ctx, uint32_t alpha, int ebwt_mode, int strand, uint64_t *dev_sent)
{
    if (!ctx || !dev_sent) return LIME_ERR_ARG;
    XchgState &S = ctx->xchg;
    if (!S.active) return fail(ctx, LIME_ERR_STATE, "lime_xchg_accumulate: no exchange set up");
    if (strand < 0 || strand >= S.n_strands) return fail(ctx, LIME_ERR_ARG, "lime_xchg_accumulate: bad strand");
    if (!ctx->loaded) return fail(ctx, LIME_ERR_STATE, "lime_xchg_accumulate: nothing loaded");
    if (!ctx->d_lcp && ctx->n_own) return fail(ctx, LIME_ERR_STATE, "lime_xchg_accumulate: loaded without lcp");
    if (ctx->sim_done) return fail(ctx, LIME_ERR_STATE, "lime_xchg_accumulate: already accumulated for this load");
    if (ebwt_mode && !ctx->d_ebwt) return fail(ctx, LIME_ERR_ARG, "lime_xchg_accumulate: EBWT mode needs the ebwt array");
    if ((uint64_t)S.per * S.world < ctx->n_reads) return fail(ctx, LIME_ERR_ARG, "lime_xchg_accumulate: read ranges do not cover the reads");
    if ((uint64_t)ctx->n_reads + ctx->n_refs >= K12_ID_LIMIT) return fail(ctx, LIME_ERR_ARG, "lime_xchg_accumulate: too many documents for the fused kernel");
    if (lime_sim_pitch(ctx->n_refs) != S.pitch) return fail(ctx, LIME_ERR_ARG, "lime_xchg_accumulate: n_refs differs from lime_xchg_alloc");
    CK(cudaSetDevice(ctx->device), "cudaSetDevice");
    int rc = reserve_score_scratch(ctx);
    if (rc) return rc;
    for (int i = 0; i < 4; i++) ctx->sim_stats[i] = 0;
    ctx->have_clusters = false;
    unsigned long long *sc = strand_scalar<unsigned long long>(ctx, strand, 0);
    CK(cudaMemsetAsync(sc, 0, sizeof(unsigned long long) * SS_COUNT, ctx->stream), "memset strand state");
    ctx->sim_done = true;
    FlatOut F;
    F.rec = S.flat.as<uint2>(); F.count = sc + SS_FLAT_COUNT; F.cap = S.flat_cap;
    rc = queue_k12_score(ctx, alpha, ebwt_mode != 0, sc, F);
    if (rc) return rc;
    XchgRoute R = {};
    R.flat = F.rec; R.flat_count = F.count; R.flat_cap = F.cap;
    R.world = (uint32_t)S.world; R.per = S.per; R.n_refs = ctx->n_refs;
    R.n_sub = S.n_sub; R.sub_rows = S.sub_rows;
    // region of this source in owner o's inbox (loopback: all owners are this GPU, the owner index picks the region)
    for (int o = 0; o < S.world; o++)
        R.dst[o] = static_cast<uint2 *>(S.inbox_map[o]) + ((uint64_t)strand * S.world + (S.loopback ? o : S.rank)) * S.inbox_cap;
    R.sub_cap = S.sub_cap; R.sent = sc + SS_SENT; R.err = sc + SS_ERR;
    {
        LaunchScope ls(ctx, KID_X_ROUTE);
        CK(launch_xchg_route(R, ctx->stream), "x_route launch");
    }
    CK(cudaMemcpyAsync(dev_sent, sc + SS_SENT, sizeof(unsigned long long) * S.world * S.n_sub, cudaMemcpyDeviceToDevice, ctx->stream), "D2D record counts");
    return LIME_OK;
}

int lime_xchg_apply(lime_ctx *ctx, int strand, const uint64_t *dev_counts_all)
{
    if (!ctx || !dev_counts_all) return LIME_ERR_ARG;
    XchgState &S = ctx->xchg;
    if (!S.active) return fail(ctx, LIME_ERR_STATE, "lime_xchg_apply: no exchange set up");
    if (strand < 0 || strand >= S.n_strands) return fail(ctx, LIME_ERR_ARG, "lime_xchg_apply: bad strand");
    CK(cudaSetDevice(ctx->device), "cudaSetDevice");
    unsigned long long *sc = strand_scalar<unsigned long long>(ctx, strand, 0);
    XchgApply A = {};
    A.src = static_cast<const uint2 *>(S.inbox_own) + (uint64_t)strand * S.world * S.inbox_cap;
    A.sub_cap = S.sub_cap;
    A.counts_all = reinterpret_cast<const unsigned long long *>(dev_counts_all);
    A.world = (uint32_t)S.world; A.rank = (uint32_t)S.rank; A.n_sub = S.n_sub;
    A.slab = S.slabs.as<uint8_t>() + (uint64_t)strand * S.rows * S.pitch; A.pitch = S.pitch;
    A.rows = (uint32_t)S.rows; A.n_refs = ctx->n_refs;
    A.applied_t = sc + SS_APPLIED_T; A.applied_n = sc + SS_APPLIED_N; A.err = sc + SS_ERR;
    for (uint32_t sub = 0; sub < S.n_sub; sub++) {             // row block after row block: the slab lines stay in L2
        A.sub = sub;
        LaunchScope ls(ctx, KID_X_APPLY);
        CK(launch_xchg_apply(A, ctx->stream), "x_apply launch");
    }
    // the slab is the matrix the choose pass reads (row 0 = first owned read)
    ctx->d_sim = A.slab; ctx->sim_rows = S.rows; ctx->sim_pitch = S.pitch;
    ctx->sim_done = true; ctx->sim_bound = false; ctx->sim_clean = false;
    return LIME_OK;
}

int lime_xchg_choose(lime_ctx *ctx, int strand, uint32_t norm, float beta, uint32_t n_own_reads)
{
    if (!ctx) return LIME_ERR_ARG;
    XchgState &S = ctx->xchg;
    if (!S.active) return fail(ctx, LIME_ERR_STATE, "lime_xchg_choose: no exchange set up");
    if (strand < 0 || strand >= S.n_strands || n_own_reads > S.rows) return fail(ctx, LIME_ERR_ARG, "lime_xchg_choose: bad argument");
    CK(cudaSetDevice(ctx->device), "cudaSetDevice");
    ctx->d_sim = S.slabs.as<uint8_t>() + (uint64_t)strand * S.rows * S.pitch; ctx->sim_rows = S.rows; ctx->sim_pitch = S.pitch;
    // every non-zero cell received a record: the inbox capacity bounds the records of the sparse rows
    const uint64_t cells = (uint64_t)n_own_reads * ctx->n_refs;
    const uint64_t cap = 1 + (uint64_t)n_own_reads + std::min<uint64_t>(cells, S.inbox_cap * (uint64_t)S.world);
    return queue_choose_deferred(ctx, strand, norm, beta, n_own_reads, cap, 2);
}

int lime_xchg_status(lime_ctx *ctx, int strand, uint64_t out[8])
{
    if (!ctx || !out) return LIME_ERR_ARG;
    XchgState &S = ctx->xchg;
    if (!S.active) return fail(ctx, LIME_ERR_STATE, "lime_xchg_status: no exchange set up");
    if (strand < 0 || strand >= S.n_strands) return fail(ctx, LIME_ERR_ARG, "lime_xchg_status: bad strand");
    CK(cudaSetDevice(ctx->device), "cudaSetDevice");
    unsigned long long *h = ctx->strand_host + (size_t)strand * SS_COUNT;
    CK(cudaMemcpyAsync(h, strand_scalar<unsigned long long>(ctx, strand, 0), sizeof(unsigned long long) * SS_COUNT,
                       cudaMemcpyDeviceToHost, ctx->stream), "D2H strand state");
    CK(cudaStreamSynchronize(ctx->stream), "lime_xchg_status");
    out[0] = h[SS_NCLUSTERS]; out[1] = h[SS_MAXLEN]; out[2] = h[SS_NUPDATES]; out[3] = h[SS_FLAT_COUNT];
    out[4] = h[SS_APPLIED_T]; out[5] = h[SS_BYTESUM]; out[6] = h[SS_NRECORDS]; out[7] = h[SS_ERR];
    return LIME_OK;
}

} // extern "C"
