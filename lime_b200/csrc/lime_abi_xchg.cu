// lime_abi_xchg.cu -- multi-GPU accumulation: similarity rows live on the GPU that owns the read (one process
// per GPU, one NVLink / NVSwitch node), updates travel there as records (kernels: k12_stream.cu MODE_FLAT,
// xchg.cu).
//
// The position range of a strand is sharded over the GPUs, the reads are owned by contiguous ranges.  Every
// GPU keeps only the rows of ITS reads (a slab per strand) and an inbox with one region per source GPU, which
// the other processes map through CUDA IPC.  Per strand and GPU:
//   lime_xchg_accumulate   context stream: fused kernel + block path append {read, genome << 8 | t} records to a
//                          local list.  Side stream: route_kernel sorts them by owner and writes each owner's run
//                          into this GPU's region of that owner's inbox (bulk stores over NVLink), publish_kernel
//                          leaves the record counts and a flag in the owner's mailbox;
//   lime_xchg_apply        side stream: wait_kernel (flags of all sources), then the records of all regions are
//                          added to the slab (local fire-and-forget adds);
//   lime_xchg_choose       side stream: clusterChoose on the slab (also: byte sum for the carry proof, rows left zero).
// There is no collective and no host synchronisation: the side-stream chain of strand s runs beside the scoring of
// strand s + 1 (two record lists, used alternately).  lime_classify joins the streams, reads the strands' state
// back and reports LIME_ERR_RETRY when a strand must be redone through the exact route (a cell above 255, a list
// too small).  Between two jobs the ranks must meet (any collective): a job rewrites the inboxes of the one before.
#include "abi_internal.cuh"
#include <algorithm>
#include <cstdlib>
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
    if (S.tail) cudaStreamSynchronize(S.tail);
    if (S.inbox_own) cudaFree(S.inbox_own);
    S.inbox_own = nullptr;
    S.slabs.release(); S.flat[0].release(); S.flat[1].release();
    auto drop = [](cudaEvent_t &e) { if (e) cudaEventDestroy(e); e = nullptr; };
    for (int k = 0; k < LIME_MAXFILES; k++) { drop(S.ev_scored[k]); drop(S.ev_strand[k]); S.strand_set[k] = false; S.epoch[k] = 0; }
    for (int k = 0; k < 2; k++) { drop(S.ev_routed[k]); S.routed_set[k] = false; }
    drop(S.ev_join);
    if (S.tail) cudaStreamDestroy(S.tail);
    S.tail = nullptr; S.tail_dirty = false;
    if (S.active) { ctx->d_sim = nullptr; ctx->sim_rows = 0; ctx->sim_clean = false; }
    S.active = false; S.world = 0; S.n_strands = 0; S.loopback = false;
}

int xchg_join(lime_ctx *ctx)
{
    XchgState &S = ctx->xchg;
    if (!S.tail || !S.tail_dirty) return LIME_OK;
    CK(cudaEventRecord(S.ev_join, S.tail), "cudaEventRecord");
    CK(cudaStreamWaitEvent(ctx->stream, S.ev_join, 0), "cudaStreamWaitEvent");
    S.tail_dirty = false;
    return LIME_OK;
}

namespace {
// launches of this scope go to the exchange's side stream
struct OnTail {
    lime_ctx *c; cudaStream_t keep;
    explicit OnTail(lime_ctx *c_) : c(c_), keep(c_->stream)
    {
        if (c->xchg.use_tail) { c->stream = c->xchg.tail; c->xchg.tail_dirty = true; }
    }
    ~OnTail() { c->stream = keep; }
};
int mark_strand(lime_ctx *ctx, int strand)          // called with the side stream current
{
    XchgState &S = ctx->xchg;
    CK(cudaEventRecord(S.ev_strand[strand], ctx->stream), "cudaEventRecord");
    S.strand_set[strand] = true;
    return LIME_OK;
}
uint64_t *mail_entry(const XchgState &S, int owner, int strand, int source)
{
    return static_cast<uint64_t *>(S.inbox_map[owner]) + ((uint64_t)strand * S.world + source) * XCHG_MAIL_STRIDE;
}
uint2 *region(const XchgState &S, void *inbox, int strand, int source)
{
    return reinterpret_cast<uint2 *>(static_cast<uint8_t *>(inbox) + S.mail_bytes) + ((uint64_t)strand * S.world + source) * S.inbox_cap;
}
} // namespace

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
    uint64_t sub_bytes = 32ull << 20;
    if (const char *e = getenv("LIME_XCHG_SUB_MB")) sub_bytes = std::max<uint64_t>(1, (uint64_t)atoll(e)) << 20;
    S.n_sub = (uint32_t)std::min<uint64_t>(XCHG_SUB_MAX, std::max<uint64_t>(1, (S.rows * S.pitch + sub_bytes - 1) / sub_bytes));
    S.sub_rows = (uint32_t)((S.rows + S.n_sub - 1) / S.n_sub);
    S.sub_cap = std::max<uint64_t>(((inbox_cap + S.n_sub - 1) / S.n_sub * 5 / 4 + 255) / 256 * 256, 256);
    S.inbox_cap = S.sub_cap * S.n_sub;
    const uint64_t slab_bytes = std::max<uint64_t>(S.rows * S.pitch * (uint64_t)n_strands, 16);
    CK(S.slabs.reserve(slab_bytes), "cudaMalloc slabs");
    CK(cudaMemsetAsync(S.slabs.p, 0, slab_bytes, ctx->stream), "memset slabs");
    CK(S.flat[0].reserve(S.flat_cap * sizeof(uint2)), "cudaMalloc update list");
    CK(S.flat[1].reserve(n_strands > 1 ? S.flat_cap * sizeof(uint2) : 16), "cudaMalloc update list");
    S.mail_bytes = ((uint64_t)n_strands * world * XCHG_MAIL_STRIDE * sizeof(uint64_t) + 4095) / 4096 * 4096;
    const uint64_t inbox_bytes = S.mail_bytes + S.inbox_cap * sizeof(uint2) * (uint64_t)world * (uint64_t)n_strands;
    CK(cudaMalloc(&S.inbox_own, inbox_bytes), "cudaMalloc inbox");          // plain cudaMalloc: IPC needs it
    CK(cudaMemsetAsync(S.inbox_own, 0, S.mail_bytes, ctx->stream), "memset mailbox");
    {
        int lo = 0, hi = 0;
        CK(cudaDeviceGetStreamPriorityRange(&lo, &hi), "cudaDeviceGetStreamPriorityRange");
        CK(cudaStreamCreateWithPriority(&S.tail, cudaStreamNonBlocking, hi), "cudaStreamCreate");
        for (int k = 0; k < LIME_MAXFILES; k++) {
            CK(cudaEventCreateWithFlags(&S.ev_scored[k], cudaEventDisableTiming), "cudaEventCreate");
            CK(cudaEventCreateWithFlags(&S.ev_strand[k], cudaEventDisableTiming), "cudaEventCreate");
        }
        for (int k = 0; k < 2; k++) CK(cudaEventCreateWithFlags(&S.ev_routed[k], cudaEventDisableTiming), "cudaEventCreate");
        CK(cudaEventCreateWithFlags(&S.ev_join, cudaEventDisableTiming), "cudaEventCreate");
    }
    // CTAs per SM the scoring kernel takes (0 = all 8).  Leaving slots to the side-stream kernels of the strand
    // before was measured (2 GPUs, cfg 2, ms per step): all 8 -> 5.96, 7 -> 6.06, 6 -> 6.30: the scoring kernel
    // slows down by 4 % per slot while route_kernel, at one CTA per SM, takes the whole scoring pass to finish.  So the
    // side stream only fills the SMs the scoring kernel drains at its end; what it buys is the absence of any
    // collective or host synchronisation between the strands (6.61 -> 5.96 ms).
    S.k12_ctas = 0;
    S.use_tail = true;
    if (const char *e = getenv("LIME_XCHG_TAIL")) S.use_tail = atoi(e) != 0;     // 0: everything in the context's stream (A/B measurements)
    if (const char *e = getenv("LIME_XCHG_K12_CTAS")) S.k12_ctas = atoi(e);
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

int lime_xchg_close(lime_ctx *ctx)
{
    if (!ctx) return LIME_ERR_ARG;
    CK(cudaSetDevice(ctx->device), "cudaSetDevice");
    { const int rc = xchg_join(ctx); if (rc) return rc; }
    CK(cudaStreamSynchronize(ctx->stream), "lime_xchg_close");
    xchg_release(ctx);
    return LIME_OK;
}

int lime_xchg_accumulate(lime_ctx *ctx, uint32_t alpha, int ebwt_mode, int strand)
{
    if (!ctx) return LIME_ERR_ARG;
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
    const int fl = strand & 1;
    // the strand's state block and its list are free once the side-stream work that used them has run
    if (S.strand_set[strand]) CK(cudaStreamWaitEvent(ctx->stream, S.ev_strand[strand], 0), "cudaStreamWaitEvent");
    if (S.routed_set[fl]) CK(cudaStreamWaitEvent(ctx->stream, S.ev_routed[fl], 0), "cudaStreamWaitEvent");
    unsigned long long *sc = strand_scalar<unsigned long long>(ctx, strand, 0);
    CK(cudaMemsetAsync(sc, 0, sizeof(unsigned long long) * SS_COUNT, ctx->stream), "memset strand state");
    ctx->sim_done = true;
    FlatOut F;
    F.rec = S.flat[fl].as<uint2>(); F.count = sc + SS_FLAT_COUNT; F.cap = S.flat_cap;
    rc = queue_k12_score(ctx, alpha, ebwt_mode != 0, sc, F, S.k12_ctas);
    if (rc) return rc;
    CK(cudaEventRecord(S.ev_scored[strand], ctx->stream), "cudaEventRecord");
    const unsigned long long epoch = ++S.epoch[strand];
    XchgRoute R = {};
    R.flat = F.rec; R.flat_count = F.count; R.flat_cap = F.cap;
    R.world = (uint32_t)S.world; R.per = S.per; R.n_refs = ctx->n_refs;
    R.n_sub = S.n_sub; R.sub_rows = S.sub_rows;
    XchgPublish Pb = {};
    Pb.sent = sc + SS_SENT; Pb.world = (uint32_t)S.world; Pb.n_sub = S.n_sub; Pb.epoch = epoch;
    // region / mailbox entry of this source at owner o (loopback: all owners are this GPU, the owner index picks them)
    for (int o = 0; o < S.world; o++) {
        const int me = S.loopback ? o : S.rank;
        R.dst[o] = region(S, S.inbox_map[o], strand, me);
        Pb.mail[o] = reinterpret_cast<unsigned long long *>(mail_entry(S, o, strand, me));
    }
    R.sub_cap = S.sub_cap; R.sent = sc + SS_SENT; R.err = sc + SS_ERR;
    OnTail on(ctx);
    CK(cudaStreamWaitEvent(ctx->stream, S.ev_scored[strand], 0), "cudaStreamWaitEvent");
    {
        LaunchScope ls(ctx, KID_X_ROUTE);
        CK(launch_xchg_route(R, ctx->stream), "x_route launch");
    }
    {
        LaunchScope ls(ctx, KID_X_PUBLISH);
        CK(launch_xchg_publish(Pb, ctx->stream), "x_publish launch");
    }
    CK(cudaEventRecord(S.ev_routed[fl], ctx->stream), "cudaEventRecord");
    S.routed_set[fl] = true;
    return mark_strand(ctx, strand);
}

int lime_xchg_apply(lime_ctx *ctx, int strand)
{
    if (!ctx) return LIME_ERR_ARG;
    XchgState &S = ctx->xchg;
    if (!S.active) return fail(ctx, LIME_ERR_STATE, "lime_xchg_apply: no exchange set up");
    if (strand < 0 || strand >= S.n_strands) return fail(ctx, LIME_ERR_ARG, "lime_xchg_apply: bad strand");
    if (!S.epoch[strand]) return fail(ctx, LIME_ERR_STATE, "lime_xchg_apply: lime_xchg_accumulate first");
    CK(cudaSetDevice(ctx->device), "cudaSetDevice");
    unsigned long long *sc = strand_scalar<unsigned long long>(ctx, strand, 0);
    const unsigned long long *mail = reinterpret_cast<const unsigned long long *>(mail_entry(S, S.rank, strand, 0));
    OnTail on(ctx);
    XchgWait W = {};
    W.mail = mail; W.world = (uint32_t)S.world; W.only = S.loopback ? S.rank : -1; W.epoch = S.epoch[strand]; W.err = sc + SS_ERR;
    {
        LaunchScope ls(ctx, KID_X_WAIT);
        CK(launch_xchg_wait(W, ctx->stream), "x_wait launch");
    }
    XchgApply A = {};
    A.src = region(S, S.inbox_own, strand, 0);
    A.sub_cap = S.sub_cap;
    A.mail = mail; A.only = W.only;
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
    return mark_strand(ctx, strand);
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
    OnTail on(ctx);
    const int rc = queue_choose_deferred(ctx, strand, norm, beta, n_own_reads, cap, 2);
    return rc ? rc : mark_strand(ctx, strand);
}

int lime_xchg_status(lime_ctx *ctx, int strand, uint64_t out[8])
{
    if (!ctx || !out) return LIME_ERR_ARG;
    XchgState &S = ctx->xchg;
    if (!S.active) return fail(ctx, LIME_ERR_STATE, "lime_xchg_status: no exchange set up");
    if (strand < 0 || strand >= S.n_strands) return fail(ctx, LIME_ERR_ARG, "lime_xchg_status: bad strand");
    CK(cudaSetDevice(ctx->device), "cudaSetDevice");
    { const int rc = xchg_join(ctx); if (rc) return rc; }
    unsigned long long *h = ctx->strand_host + (size_t)strand * SS_COUNT;
    CK(cudaMemcpyAsync(h, strand_scalar<unsigned long long>(ctx, strand, 0), sizeof(unsigned long long) * SS_COUNT,
                       cudaMemcpyDeviceToHost, ctx->stream), "D2H strand state");
    CK(cudaStreamSynchronize(ctx->stream), "lime_xchg_status");
    out[0] = h[SS_NCLUSTERS]; out[1] = h[SS_MAXLEN]; out[2] = h[SS_NUPDATES]; out[3] = h[SS_FLAT_COUNT];
    out[4] = h[SS_APPLIED_T]; out[5] = h[SS_BYTESUM]; out[6] = h[SS_NRECORDS]; out[7] = h[SS_ERR];
    return LIME_OK;
}

} // extern "C"
