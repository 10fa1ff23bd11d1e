// lime_abi_peer.cu -- multi-GPU accumulation over peer memory (one process per GPU, one NVSwitch node).
//
// The position range of a strand is sharded over the GPUs, the reads are owned by contiguous ranges.
// Instead of a partial matrix over all reads per GPU followed by an exchange, every GPU keeps only the
// rows of ITS reads (a "slab", one per strand) and maps the slabs of all other GPUs through CUDA IPC;
// the fused kernel then adds each (read, genome, t) straight into the owner's slab -- a fire-and-forget
// `red` that NVLink / NVSwitch carries to the owner's L2.  Compute and exchange are one kernel.
// Carries out of a uint8 cell are ruled out as on one GPU: the byte sum of all slabs must equal the
// total of all t added by all GPUs (the caller reduces both numbers).
#include "abi_internal.cuh"
#include <algorithm>
#include <cstring>

static_assert(sizeof(cudaIpcMemHandle_t) == LIME_IPC_HANDLE_BYTES, "IPC handle size");
static_assert(PEER_MAX == LIME_PEER_MAX, "peer limit");

lime::PeerTable peer_table(const lime_ctx *ctx, int slab)
{
    const PeerState &S = ctx->peer;
    PeerTable T;
    memset(&T, 0, sizeof(T));
    for (int k = 0; k < S.world; k++)
        T.base[k] = static_cast<uint8_t *>(S.mapped[k]) + (uint64_t)slab * S.rows * S.pitch;
    T.per = S.per;
    T.magic = (uint32_t)std::min<uint64_t>((1ull << 32) / std::max<uint32_t>(S.per, 1u), 0xffffffffull);
    return T;
}

void peer_release(lime_ctx *ctx)
{
    PeerState &S = ctx->peer;
    for (int k = 0; k < S.world; k++)
        if (S.mapped[k] && S.mapped[k] != S.own) cudaIpcCloseMemHandle(S.mapped[k]);
    for (int k = 0; k < PEER_MAX; k++) S.mapped[k] = nullptr;
    if (S.own) cudaFree(S.own);
    S.own = nullptr;
    S.tables.release();
    if (S.active) { ctx->d_sim = nullptr; ctx->sim_rows = 0; }
    S.active = false; S.world = 0; S.n_slabs = 0;
}

// ---- instrumentation: rate of random adds into the slabs -----------------------------------------------
// mode 0: fire-and-forget add of 0 (red.or), 1: add with returned value (atom), 2: plain 32-bit store,
// 3: fire-and-forget add of a non-zero value (red.add, what the fused kernel issues)
__global__ void peer_probe_kernel(PeerTable T, uint64_t pitch, uint32_t span_rows, uint32_t n_refs, int first, int count,
                                  uint64_t n_adds, int mode, unsigned *sink)
{
    unsigned acc = 0;
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n_adds; i += (uint64_t)gridDim.x * blockDim.x) {
        uint64_t h = (i + 1) * 0x9E3779B97F4A7C15ull;
        h ^= h >> 29; h *= 0xBF58476D1CE4E5B9ull; h ^= h >> 32;
        const uint32_t owner = (uint32_t)first + (uint32_t)(h % (uint32_t)count);
        const uint32_t row = (uint32_t)((h >> 8) % span_rows);
        const uint32_t g = (uint32_t)((h >> 40) % n_refs);
        unsigned *w = reinterpret_cast<unsigned *>(T.base[owner] + (size_t)row * pitch) + (g >> 2);
        if (mode == 0) atomicAdd(w, 0u);
        else if (mode == 1) acc += atomicAdd(w, 0u);
        else if (mode == 3) atomicAdd(w, 1u << ((g & 3u) * 8u));
        else *w = 0u;
    }
    if (acc == 0xdeadbeefu) *sink = acc;
}

extern "C" {

int lime_peer_probe(lime_ctx *ctx, int slab, int first_owner, int n_owners, uint32_t span_rows, uint64_t n_adds, int mode, double *ms)
{
    if (!ctx || !ms) return LIME_ERR_ARG;
    PeerState &S = ctx->peer;
    if (!S.active) return fail(ctx, LIME_ERR_STATE, "lime_peer_probe: no peer slabs");
    if (slab < 0 || slab >= S.n_slabs || first_owner < 0 || n_owners < 1 || first_owner + n_owners > S.world ||
        span_rows == 0 || span_rows > S.rows)
        return fail(ctx, LIME_ERR_ARG, "lime_peer_probe: bad argument");
    CK(cudaSetDevice(ctx->device), "cudaSetDevice");
    cudaEvent_t a, b;
    CK(cudaEventCreate(&a), "event"); CK(cudaEventCreate(&b), "event");
    const PeerTable T = peer_table(ctx, slab);
    const uint32_t n_refs = (uint32_t)std::max<uint64_t>(S.pitch, 4);
    CK(cudaEventRecord(a, ctx->stream), "event");
    peer_probe_kernel<<<148 * 8, 256, 0, ctx->stream>>>(T, S.pitch, span_rows, n_refs, first_owner, n_owners, n_adds, mode,
                                                        scalar<unsigned int>(ctx, SC_TILE_COUNTER));
    CK(cudaGetLastError(), "peer_probe launch");
    CK(cudaEventRecord(b, ctx->stream), "event");
    CK(cudaEventSynchronize(b), "peer_probe");
    float f = 0;
    CK(cudaEventElapsedTime(&f, a, b), "event");
    cudaEventDestroy(a); cudaEventDestroy(b);
    *ms = f;
    return LIME_OK;
}

int lime_peer_alloc(lime_ctx *ctx, uint64_t rows, uint32_t n_refs, int n_slabs, uint8_t *ipc_handle)
{
    if (!ctx || !ipc_handle || n_slabs < 1 || n_slabs > LIME_MAXFILES) return LIME_ERR_ARG;
    CK(cudaSetDevice(ctx->device), "cudaSetDevice");
    CK(cudaStreamSynchronize(ctx->stream), "lime_peer_alloc");
    peer_release(ctx);
    PeerState &S = ctx->peer;
    S.rows = rows; S.pitch = lime_sim_pitch(n_refs); S.n_slabs = n_slabs;
    const uint64_t bytes = std::max<uint64_t>(S.rows * S.pitch * (uint64_t)n_slabs, 16);
    CK(cudaMalloc(&S.own, bytes), "cudaMalloc slabs");         // plain cudaMalloc: IPC needs it
    CK(cudaMemsetAsync(S.own, 0, bytes, ctx->stream), "memset slabs");
    cudaIpcMemHandle_t h;
    CK(cudaIpcGetMemHandle(&h, S.own), "cudaIpcGetMemHandle");
    memcpy(ipc_handle, &h, sizeof(h));
    CK(cudaStreamSynchronize(ctx->stream), "lime_peer_alloc");
    return LIME_OK;
}

int lime_peer_open(lime_ctx *ctx, int world, int rank, const uint8_t *ipc_handles, uint32_t reads_per_rank)
{
    if (!ctx || !ipc_handles || world < 1 || world > LIME_PEER_MAX || rank < 0 || rank >= world) return LIME_ERR_ARG;
    PeerState &S = ctx->peer;
    if (!S.own) return fail(ctx, LIME_ERR_STATE, "lime_peer_open: lime_peer_alloc first");
    if (reads_per_rank == 0 || reads_per_rank > S.rows) return fail(ctx, LIME_ERR_ARG, "lime_peer_open: slab smaller than a read range");
    CK(cudaSetDevice(ctx->device), "cudaSetDevice");
    S.world = world; S.rank = rank; S.per = reads_per_rank;
    for (int k = 0; k < world; k++) {
        if (k == rank) { S.mapped[k] = S.own; continue; }
        cudaIpcMemHandle_t h;
        memcpy(&h, ipc_handles + (size_t)k * LIME_IPC_HANDLE_BYTES, sizeof(h));
        CK(cudaIpcOpenMemHandle(&S.mapped[k], h, cudaIpcMemLazyEnablePeerAccess), "cudaIpcOpenMemHandle");
    }
    CK(S.tables.reserve(sizeof(PeerTable) * S.n_slabs), "cudaMalloc peer tables");
    for (int s = 0; s < S.n_slabs; s++) {
        const PeerTable T = peer_table(ctx, s);
        CK(cudaMemcpyAsync(S.tables.as<PeerTable>() + s, &T, sizeof(T), cudaMemcpyHostToDevice, ctx->stream), "H2D peer table");
        CK(cudaStreamSynchronize(ctx->stream), "H2D peer table");
    }
    S.active = true;
    ctx->sim_bound = false; ctx->d_sim = nullptr; ctx->sim_rows = 0; ctx->sim_pitch = S.pitch;
    return LIME_OK;
}

int lime_peer_close(lime_ctx *ctx)
{
    if (!ctx) return LIME_ERR_ARG;
    CK(cudaSetDevice(ctx->device), "cudaSetDevice");
    CK(cudaStreamSynchronize(ctx->stream), "lime_peer_close");
    peer_release(ctx);
    return LIME_OK;
}

int lime_peer_zero(lime_ctx *ctx)
{
    if (!ctx) return LIME_ERR_ARG;
    PeerState &S = ctx->peer;
    if (!S.active) return fail(ctx, LIME_ERR_STATE, "lime_peer_zero: no peer slabs");
    CK(cudaSetDevice(ctx->device), "cudaSetDevice");
    if (S.rows * S.pitch)
        CK(cudaMemsetAsync(S.own, 0, S.rows * S.pitch * (uint64_t)S.n_slabs, ctx->stream), "memset slabs");
    return LIME_OK;
}

int lime_peer_accumulate(lime_ctx *ctx, uint32_t alpha, int ebwt_mode, int slab, uint64_t out[4])
{
    if (!ctx) return LIME_ERR_ARG;
    PeerState &S = ctx->peer;
    if (!S.active) return fail(ctx, LIME_ERR_STATE, "lime_peer_accumulate: no peer slabs");
    if (slab < 0 || slab >= S.n_slabs) return fail(ctx, LIME_ERR_ARG, "lime_peer_accumulate: bad slab");
    if (!ctx->loaded) return fail(ctx, LIME_ERR_STATE, "lime_peer_accumulate: nothing loaded");
    if (!ctx->d_lcp && ctx->n_own) return fail(ctx, LIME_ERR_STATE, "lime_peer_accumulate: loaded without lcp");
    if (ctx->sim_done) return fail(ctx, LIME_ERR_STATE, "lime_peer_accumulate: already accumulated for this load");
    if (ebwt_mode && !ctx->d_ebwt) return fail(ctx, LIME_ERR_ARG, "lime_peer_accumulate: EBWT mode needs the ebwt array");
    if ((uint64_t)S.per * S.world < ctx->n_reads) return fail(ctx, LIME_ERR_ARG, "lime_peer_accumulate: read ranges do not cover the reads");
    CK(cudaSetDevice(ctx->device), "cudaSetDevice");
    const uint64_t n_total = ctx->n_own + ctx->n_halo;
    for (int i = 0; i < 4; i++) ctx->sim_stats[i] = 0;
    ctx->have_clusters = false;
    CK(ctx->long_recs.reserve(k12_long_cap(n_total) * sizeof(lime_cluster)), "cudaMalloc long-cluster list");
    CK(ctx->k2_scratch.reserve(k2_block_scratch_bytes() * (uint64_t)k2_block_grid()), "cudaMalloc block-path scratch");
    unsigned long long sc[SC_COUNT];
    ctx->sim_done = true;
    const int rc = run_k12_pass(ctx, alpha, ebwt_mode != 0, false, false, 0, 0, sc, slab);
    if (rc) return rc;
    ctx->n_clusters = sc[SC_NCLUSTERS]; ctx->max_len = sc[SC_MAXLEN]; ctx->alpha = alpha;
    ctx->sim_stats[1] = sc[SC_LARGE_COUNT];
    ctx->sim_stats[0] = sc[SC_NCLUSTERS] - sc[SC_LARGE_COUNT];
    ctx->sim_stats[2] = sc[SC_NUPDATES];
    if (out) { out[0] = sc[SC_NCLUSTERS]; out[1] = sc[SC_MAXLEN]; out[2] = sc[SC_NUPDATES]; out[3] = sc[SC_SUMT]; }
    return LIME_OK;
}

int lime_peer_select(lime_ctx *ctx, int slab, uint64_t total_updates, uint64_t *byte_sum)
{
    if (!ctx) return LIME_ERR_ARG;
    PeerState &S = ctx->peer;
    if (!S.active) return fail(ctx, LIME_ERR_STATE, "lime_peer_select: no peer slabs");
    if (slab < 0 || slab >= S.n_slabs) return fail(ctx, LIME_ERR_ARG, "lime_peer_select: bad slab");
    CK(cudaSetDevice(ctx->device), "cudaSetDevice");
    ctx->d_sim = static_cast<uint8_t *>(S.own) + (uint64_t)slab * S.rows * S.pitch;
    ctx->sim_rows = S.rows; ctx->sim_pitch = S.pitch;
    ctx->sim_done = true; ctx->sim_bound = false;
    ctx->sim_stats[2] = total_updates;         // bounds the records of lime_choose
    if (byte_sum) {
        // asynchronous: the value is in device memory when the stream reaches this point
        unsigned long long *d = reinterpret_cast<unsigned long long *>(byte_sum);
        CK(cudaMemsetAsync(d, 0, 8, ctx->stream), "memset byte sum");
        if (S.rows * S.pitch) {
            LaunchScope ls(ctx, KID_K12_CHECK);
            CK(launch_bytesum(ctx->d_sim, S.rows * S.pitch, d, ctx->stream), "k12_check launch");
        }
    }
    return LIME_OK;
}

int lime_peer_zero_slab(lime_ctx *ctx, int slab)
{
    if (!ctx) return LIME_ERR_ARG;
    PeerState &S = ctx->peer;
    if (!S.active) return fail(ctx, LIME_ERR_STATE, "lime_peer_zero_slab: no peer slabs");
    if (slab < 0 || slab >= S.n_slabs) return fail(ctx, LIME_ERR_ARG, "lime_peer_zero_slab: bad slab");
    CK(cudaSetDevice(ctx->device), "cudaSetDevice");
    if (S.rows * S.pitch)
        CK(cudaMemsetAsync(static_cast<uint8_t *>(S.own) + (uint64_t)slab * S.rows * S.pitch, 0, S.rows * S.pitch, ctx->stream),
           "memset slab");
    return LIME_OK;
}

} // extern "C"
