// lime_abi.cu -- the C ABI of include/lime_cuda.h: context, residency, step drivers.
// No CPU fallback: every entry point needs a CUDA device; errors come back as LIME_ERR_* codes.
#include "abi_internal.cuh"
#include <cstdio>
#include <cstring>
#include <algorithm>

static std::string g_init_error;

extern "C" {

const char *lime_version(void) { return "lime-b200 0.2 (sm_100a)"; }

const char *lime_last_error(const lime_ctx *ctx) { return ctx ? ctx->err.c_str() : g_init_error.c_str(); }

int lime_init(lime_ctx **out, int device)
{
    if (!out) return LIME_ERR_ARG;
    *out = nullptr;
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess || n == 0) {
        g_init_error = std::string("no CUDA device: ") + cudaGetErrorString(e);
        return LIME_ERR_CUDA;
    }
    if (device < 0 || device >= n) { g_init_error = "device ordinal out of range"; return LIME_ERR_ARG; }
    e = cudaSetDevice(device);
    if (e != cudaSuccess) { g_init_error = cudaGetErrorString(e); return LIME_ERR_CUDA; }
    lime_ctx *ctx = new lime_ctx();
    ctx->device = device;
    e = cudaStreamCreateWithFlags(&ctx->own_stream, cudaStreamNonBlocking);
    if (e == cudaSuccess) e = ctx->scalars.reserve(sizeof(unsigned long long) * SC_COUNT);
    if (e == cudaSuccess) e = k12_policies(ctx->l2_policy);
    if (e != cudaSuccess) { g_init_error = cudaGetErrorString(e); delete ctx; return LIME_ERR_CUDA; }
    ctx->stream = ctx->own_stream;
    *out = ctx;
    return LIME_OK;
}

void lime_destroy(lime_ctx *ctx)
{
    if (!ctx) return;
    cudaSetDevice(ctx->device);
    cudaStreamSynchronize(ctx->stream);
    for (auto &p : ctx->prof_events) { cudaEventDestroy(p.a); cudaEventDestroy(p.b); }
    xchg_release(ctx);
    load_release(ctx);
    if (ctx->strand_host) cudaFreeHost(ctx->strand_host);
    DevBuf *bufs[] = {&ctx->in_lcp, &ctx->in_da, &ctx->in_ebwt, &ctx->clusters, &ctx->tile_state,
                      &ctx->scalars, &ctx->strand_state, &ctx->sim_own, &ctx->sim32, &ctx->large_list, &ctx->long_recs, &ctx->k2_scratch,
                      &ctx->row_counts, &ctx->cls_out, &ctx->tax_tables};
    for (DevBuf *b : bufs) b->release();
    for (auto &s : ctx->slots) { s.bin.release(); s.pos.release(); }
    if (ctx->own_stream) cudaStreamDestroy(ctx->own_stream);
    delete ctx;
}

int lime_set_stream(lime_ctx *ctx, void *cuda_stream)
{
    if (!ctx) return LIME_ERR_ARG;
    cudaStreamSynchronize(ctx->stream);
    ctx->stream = cuda_stream ? reinterpret_cast<cudaStream_t>(cuda_stream) : ctx->own_stream;
    return LIME_OK;
}

void *lime_pinned_alloc(uint64_t bytes)
{
    void *p = nullptr;
    if (cudaHostAlloc(&p, bytes ? bytes : 1, cudaHostAllocDefault) != cudaSuccess) { cudaGetLastError(); return nullptr; }
    return p;
}
void lime_pinned_free(void *p) { if (p) cudaFreeHost(p); }

uint64_t lime_sim_pitch(uint32_t n_refs) { return ((uint64_t)n_refs + 15) / 16 * 16; }

int lime_sim_bind(lime_ctx *ctx, void *dev_matrix, uint64_t rows, uint64_t pitch)
{
    if (!ctx) return LIME_ERR_ARG;
    if (!dev_matrix) { ctx->sim_bound = false; ctx->d_sim = nullptr; return LIME_OK; }
    if (pitch % 16 != 0 || (reinterpret_cast<uintptr_t>(dev_matrix) & 15))
        return fail(ctx, LIME_ERR_ARG, "lime_sim_bind: matrix and pitch must be 16-byte aligned");
    ctx->sim_bound = true;
    ctx->d_sim = static_cast<uint8_t *>(dev_matrix);
    ctx->sim_rows = rows; ctx->sim_pitch = pitch;
    return LIME_OK;
}

int lime_load(lime_ctx *ctx, const uint32_t *lcp, const uint32_t *da, const uint8_t *ebwt,
              uint64_t n_own, uint64_t n_halo, uint64_t pos_base, int is_last,
              uint32_t n_reads, uint32_t n_refs, int mem)
{
    if (!ctx) return LIME_ERR_ARG;
    const uint64_t n = n_own + n_halo;
    // lcp may be NULL when only step 2 runs from a cluster list (ClusterBWT_DA never reads .lcp)
    if (n > 0 && !da) return fail(ctx, LIME_ERR_ARG, "lime_load: da is NULL");
    if ((uint64_t)n_reads + n_refs > 0xfffffffeull) return fail(ctx, LIME_ERR_ARG, "lime_load: too many documents");
    CK(cudaSetDevice(ctx->device), "cudaSetDevice");
    ctx->loaded = false; ctx->have_clusters = false; ctx->sim_done = false;
    if (mem == LIME_MEM_DEVICE) {
        if ((reinterpret_cast<uintptr_t>(lcp) & 15) || (reinterpret_cast<uintptr_t>(da) & 15) ||
            (reinterpret_cast<uintptr_t>(ebwt) & 3))
            return fail(ctx, LIME_ERR_ARG, "lime_load: device lcp/da must be 16-byte aligned, ebwt 4-byte aligned");
        ctx->d_lcp = lcp; ctx->d_da = da; ctx->d_ebwt = ebwt;
        ctx->lcp8 = false; ctx->lcp_alpha = 0;
    } else {
        // narrowed lcp transfer (lime_abi_load.cu): one bit per position when the caller has named alpha
        // (lime_set_alpha), one byte when it has promised alpha <= 255 (lime_set_alpha_limit)
        const bool big = lcp && n >= (1ull << 16);
        const uint32_t bits_alpha = big && ctx->alpha_exact >= 1 && ctx->alpha_exact <= 255 ? ctx->alpha_exact : 0;
        const bool narrow = bits_alpha || (big && ctx->alpha_limit >= 1 && ctx->alpha_limit <= 255);
        if ((lcp && !narrow) || n == 0) CK(ctx->in_lcp.reserve(std::max<uint64_t>(n, 4) * 4), "cudaMalloc lcp");
        CK(ctx->in_da.reserve(std::max<uint64_t>(n, 4) * 4), "cudaMalloc da");
        if (ebwt) CK(ctx->in_ebwt.reserve(std::max<uint64_t>(n, 16)), "cudaMalloc ebwt");
        if (narrow) {
            // with a named alpha da crosses as 3 bytes per position while every document id fits (LIME_PACK_DA=0: as it is)
            static const bool pack_da = !(getenv("LIME_PACK_DA") && atoi(getenv("LIME_PACK_DA")) == 0);
            const bool da3 = bits_alpha && pack_da && (uint64_t)n_reads + n_refs <= (1ull << 24);
            ctx->da3_loaded = da3;
            const int rc = load_host_narrow(ctx, lcp, da, ebwt, n, bits_alpha, da3);
            if (rc != LIME_OK) return rc;
        } else if (n) {
            if (lcp) CK(cudaMemcpyAsync(ctx->in_lcp.p, lcp, n * 4, cudaMemcpyHostToDevice, ctx->stream), "H2D lcp");
            CK(cudaMemcpyAsync(ctx->in_da.p, da, n * 4, cudaMemcpyHostToDevice, ctx->stream), "H2D da");
            if (ebwt) CK(cudaMemcpyAsync(ctx->in_ebwt.p, ebwt, n, cudaMemcpyHostToDevice, ctx->stream), "H2D ebwt");
        }
        // a narrowed load leaves min(lcp, 255) as bytes on the device; the kernels read those
        ctx->lcp8 = narrow; ctx->lcp_alpha = bits_alpha;
        ctx->d_lcp = narrow ? ctx->in_lcp8.p : ((lcp || n == 0) ? ctx->in_lcp.p : nullptr);
        ctx->d_da = ctx->in_da.as<uint32_t>();
        ctx->d_ebwt = ebwt ? ctx->in_ebwt.as<uint8_t>() : nullptr;
    }
    ctx->n_own = n_own; ctx->n_halo = n_halo; ctx->pos_base = pos_base; ctx->is_last = is_last;
    ctx->n_reads = n_reads; ctx->n_refs = n_refs;
    // similarity accumulator: dense uint8 [n_reads][pitch], zeroed
    if (ctx->xchg.active) {
        // exchange mode: the rows of the owned reads live in the slabs of lime_xchg_alloc
        ctx->loaded = true;
        return LIME_OK;
    }
    if (!ctx->sim_bound) {
        ctx->sim_pitch = lime_sim_pitch(n_refs);
        ctx->sim_rows = n_reads;
        CK(ctx->sim_own.reserve(std::max<uint64_t>(ctx->sim_rows * ctx->sim_pitch, 16)), "cudaMalloc similarity matrix");
        ctx->d_sim = ctx->sim_own.as<uint8_t>();
    } else if (ctx->sim_rows < n_reads || ctx->sim_pitch < n_refs) {
        return fail(ctx, LIME_ERR_ARG, "lime_load: bound similarity matrix is too small");
    }
    // the choose pass of lime_strand_run leaves the matrix zero: nothing to clear then
    const bool clean = ctx->sim_clean && ctx->sim_clean_ptr == ctx->d_sim && ctx->sim_clean_bytes == ctx->sim_rows * ctx->sim_pitch;
    if (ctx->sim_rows * ctx->sim_pitch && !clean)
        CK(cudaMemsetAsync(ctx->d_sim, 0, ctx->sim_rows * ctx->sim_pitch, ctx->stream), "memset similarity matrix");
    ctx->sim_clean = true; ctx->sim_clean_ptr = ctx->d_sim; ctx->sim_clean_bytes = ctx->sim_rows * ctx->sim_pitch;
    ctx->loaded = true;
    return LIME_OK;
}

int lime_cluster_detect(lime_ctx *ctx, uint32_t alpha, uint64_t *n_clusters, uint64_t *max_len)
{
    if (!ctx) return LIME_ERR_ARG;
    if (!ctx->loaded) return fail(ctx, LIME_ERR_STATE, "lime_cluster_detect: nothing loaded");
    if (!ctx->d_lcp && ctx->n_own) return fail(ctx, LIME_ERR_STATE, "lime_cluster_detect: loaded without lcp");
    unsigned long long host_sc[SC_COUNT];
    const int rc = detect_into_ctx(ctx, alpha, host_sc);
    if (rc) return rc;
    ctx->max_len = host_sc[SC_MAXLEN];
    ctx->have_clusters = true;
    if (n_clusters) *n_clusters = ctx->n_clusters;
    if (max_len) *max_len = ctx->max_len;
    return LIME_OK;
}

int lime_clusters_copy(lime_ctx *ctx, lime_cluster *out, uint64_t cap)
{
    if (!ctx || (!out && cap)) return LIME_ERR_ARG;
    if (!ctx->have_clusters) return fail(ctx, LIME_ERR_STATE, "lime_clusters_copy: no clusters");
    const uint64_t n = std::min(cap, ctx->n_clusters);
    if (n) {
        CK(cudaMemcpyAsync(out, ctx->clusters.p, n * sizeof(lime_cluster), cudaMemcpyDeviceToHost, ctx->stream), "D2H clusters");
        CK(cudaStreamSynchronize(ctx->stream), "D2H clusters");
    }
    return LIME_OK;
}

int lime_clusters_set(lime_ctx *ctx, const lime_cluster *cl, uint64_t n, uint64_t max_len)
{
    if (!ctx || (!cl && n)) return LIME_ERR_ARG;
    if (!ctx->loaded) return fail(ctx, LIME_ERR_STATE, "lime_clusters_set: nothing loaded");
    CK(cudaSetDevice(ctx->device), "cudaSetDevice");
    CK(ctx->clusters.reserve(std::max<uint64_t>(n, 1) * sizeof(lime_cluster)), "cudaMalloc clusters");
    ctx->cl_cap = ctx->clusters.cap / sizeof(lime_cluster);
    const uint64_t lo = ctx->pos_base, hi = ctx->pos_base + ctx->n_own + ctx->n_halo;
    uint64_t mx = 0;
    for (uint64_t i = 0; i < n; i++) {
        if (cl[i].pStart < lo || cl[i].pStart > hi || cl[i].len > hi - cl[i].pStart)
            return fail(ctx, LIME_ERR_ARG, "lime_clusters_set: cluster outside the resident range");
        mx = std::max(mx, cl[i].len);
    }
    if (n) {
        CK(cudaMemcpyAsync(ctx->clusters.p, cl, n * sizeof(lime_cluster), cudaMemcpyHostToDevice, ctx->stream), "H2D clusters");
        CK(cudaStreamSynchronize(ctx->stream), "H2D clusters");
    }
    ctx->n_clusters = n;
    ctx->max_len = std::max(mx, max_len);
    ctx->have_clusters = true;
    return LIME_OK;
}

} // extern "C"

// the detection passes (count, scan, emit): cluster records of this shard into ctx->clusters, ascending pStart;
// sets ctx->n_clusters / alpha, returns the scalars
int check_alpha(lime_ctx *ctx, uint32_t alpha)
{
    if (ctx->lcp8 && ctx->lcp_alpha && alpha != ctx->lcp_alpha)
        return fail(ctx, LIME_ERR_STATE, "lcp was loaded as the bit lcp < alpha of lime_set_alpha: another alpha needs a reload");
    if (ctx->lcp8 && alpha > 255)
        return fail(ctx, LIME_ERR_STATE, "lcp was loaded narrowed (lime_set_alpha_limit <= 255), alpha > 255 needs a reload");
    return LIME_OK;
}

int detect_into_ctx(lime_ctx *ctx, uint32_t alpha, unsigned long long host_sc[SC_COUNT])
{
    { const int rc = check_alpha(ctx, alpha); if (rc) return rc; }
    CK(cudaSetDevice(ctx->device), "cudaSetDevice");
    const uint64_t n_total = ctx->n_own + ctx->n_halo;
    const uint64_t tiles = k12_num_tiles(ctx->n_own);
    CK(ctx->tile_state.reserve(std::max<uint64_t>(tiles, 1) * 8), "cudaMalloc tile state");
    // capacity guess: mixed clusters are a few percent of the positions; retry once when short
    uint64_t cap = std::max<uint64_t>(ctx->cl_cap, std::max<uint64_t>(1024, n_total / 8));
    for (int attempt = 0; attempt < 2; attempt++) {
        CK(ctx->clusters.reserve(cap * sizeof(lime_cluster)), "cudaMalloc clusters");
        ctx->cl_cap = ctx->clusters.cap / sizeof(lime_cluster);
        CK(cudaMemsetAsync(ctx->scalars.p, 0, sizeof(unsigned long long) * SC_COUNT, ctx->stream), "memset scalars");
        CK(cudaMemsetAsync(ctx->tile_state.p, 0, std::max<uint64_t>(tiles, 1) * 8, ctx->stream), "memset tile state");
        K12Args a = {};
        a.lcp = ctx->d_lcp; a.lcp8 = ctx->lcp8 ? 1 : 0; a.da = ctx->d_da; a.ebwt = nullptr;
        a.n_own = ctx->n_own; a.n_total = n_total; a.pos_base = ctx->pos_base; a.is_last = ctx->is_last;
        a.alpha = alpha; a.n_reads = ctx->n_reads; a.n_refs = ctx->n_refs;
        a.out = ctx->clusters.as<lime_cluster>(); a.out_cap = ctx->cl_cap;
        a.tile_state = ctx->tile_state.as<unsigned long long>();
        a.n_clusters = scalar<unsigned long long>(ctx, SC_NCLUSTERS);
        a.max_len = scalar<unsigned long long>(ctx, SC_MAXLEN);
        a.err = scalar<unsigned int>(ctx, SC_ERR);
        a.pol_evict_first = ctx->l2_policy[0]; a.pol_evict_last = ctx->l2_policy[1];
        if (tiles) {
            LaunchScope ls(ctx, KID_K1_DETECT);
            CK(launch_k12(a, K12_DETECT, ctx->stream), "k1_detect launch");
        }
        CK(cudaMemcpyAsync(host_sc, ctx->scalars.p, sizeof(unsigned long long) * SC_COUNT, cudaMemcpyDeviceToHost, ctx->stream), "D2H scalars");
        CK(cudaStreamSynchronize(ctx->stream), "k1_detect");
        if ((unsigned)host_sc[SC_ERR] & 1u)
            return fail(ctx, LIME_ERR_HALO, "lime_cluster_detect: a cluster runs past the halo of this shard");
        if (host_sc[SC_NCLUSTERS] <= ctx->cl_cap) break;
        cap = host_sc[SC_NCLUSTERS];
    }
    ctx->n_clusters = host_sc[SC_NCLUSTERS];
    ctx->alpha = alpha;
    return LIME_OK;
}

// the remaining entry points live in lime_abi_steps.cu
