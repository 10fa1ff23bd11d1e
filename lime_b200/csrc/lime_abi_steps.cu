// lime_abi_steps.cu -- ABI entry points for step 2 (similarity, choose) and instrumentation.
#include "abi_internal.cuh"
#include <algorithm>
#include <cstdlib>
#include <vector>

static int run_k2_pass(lime_ctx *ctx, bool ebwt, bool exact32, uint32_t lo, uint32_t hi)
{
    K2Args a;
    a.cl = ctx->clusters.as<lime_cluster>(); a.n_cl = ctx->n_clusters;
    a.da = ctx->d_da; a.ebwt = ebwt ? ctx->d_ebwt : nullptr;
    a.pos_base = ctx->pos_base; a.n_reads = ctx->n_reads; a.n_refs = ctx->n_refs;
    a.sim8 = ctx->d_sim; a.pitch8 = ctx->sim_pitch;
    a.sim32 = ctx->sim32.as<uint32_t>(); a.lo = lo; a.hi = hi;
    a.carry = scalar<unsigned int>(ctx, SC_CARRY); a.sum_t = nullptr;
    a.large_count = scalar<unsigned long long>(ctx, SC_LARGE_COUNT);
    a.large_next = scalar<unsigned long long>(ctx, SC_LARGE_NEXT);
    a.n_updates = scalar<unsigned long long>(ctx, SC_NUPDATES);
    a.large_list = ctx->large_list.as<unsigned long long>(); a.large_cap = ctx->large_list.cap / 8;
    a.long_recs = nullptr;
    a.scratch = ctx->k2_scratch.as<uint8_t>(); a.scratch_per_cta = k2_block_scratch_bytes();
    CK(cudaMemsetAsync(ctx->scalars.p, 0, sizeof(unsigned long long) * SC_COUNT, ctx->stream), "memset scalars");
    {
        LaunchScope ls(ctx, KID_K2_WARP);
        CK(launch_k2_warp(a, exact32, ctx->stream), "k2_warp launch");
    }
    {
        LaunchScope ls(ctx, KID_K2_BLOCK);
        CK(launch_k2_block(a, exact32, ctx->stream, k2_block_grid()), "k2_block launch");
    }
    return LIME_OK;
}

// one fused pass (steps 1 + 2a) plus the block path over the clusters it set aside;
// peer_slab >= 0: the adds go to the owners' slabs of that index (lime_abi_peer.cu) and the carry
// check is left to the caller (it needs every GPU's adds to have landed)
int run_k12_pass(lime_ctx *ctx, uint32_t alpha, bool ebwt, bool emit, bool exact32, uint32_t lo, uint32_t hi,
                 unsigned long long sc[SC_COUNT], int peer_slab)
{
    if (ctx->lcp_clamped && alpha > 255)
        return fail(ctx, LIME_ERR_STATE, "lcp was loaded narrowed (lime_set_alpha_limit <= 255), alpha > 255 needs a reload");
    const uint64_t n_total = ctx->n_own + ctx->n_halo;
    K12Args a;
    a.lcp = ctx->d_lcp; a.da = ctx->d_da; a.ebwt = ebwt ? ctx->d_ebwt : nullptr;
    a.n_own = ctx->n_own; a.n_total = n_total; a.pos_base = ctx->pos_base; a.is_last = ctx->is_last;
    a.alpha = alpha; a.n_reads = ctx->n_reads; a.n_refs = ctx->n_refs;
    a.out = ctx->clusters.as<lime_cluster>(); a.out_cap = emit ? ctx->cl_cap : 0;
    a.tile_state = ctx->tile_state.as<unsigned long long>();
    a.n_clusters = scalar<unsigned long long>(ctx, SC_NCLUSTERS);
    a.max_len = scalar<unsigned long long>(ctx, SC_MAXLEN);
    a.err = scalar<unsigned int>(ctx, SC_ERR);
    a.long_list = ctx->long_recs.as<lime_cluster>();
    a.long_count = scalar<unsigned long long>(ctx, SC_LARGE_COUNT);
    a.long_cap = ctx->long_recs.cap / sizeof(lime_cluster);
    a.sim8 = ctx->d_sim; a.pitch8 = ctx->sim_pitch;
    a.sim32 = ctx->sim32.as<uint32_t>(); a.lo = lo; a.hi = hi;
    a.carry = scalar<unsigned int>(ctx, SC_CARRY);
    a.n_updates = scalar<unsigned long long>(ctx, SC_NUPDATES);
    { const char *e = getenv("LIME_DEBUG"); a.dbg = e ? (unsigned)atoi(e) : 0u; }
    // fire-and-forget adds; carries are ruled out after the pass by the byte sum (see below)
    a.sum_t = !exact32 ? scalar<unsigned long long>(ctx, SC_SUMT) : nullptr;
    const bool peer = peer_slab >= 0;
    if (peer) {
        a.peer_on = 1; a.peer_self = (uint32_t)ctx->peer.rank; a.peer_world = (uint32_t)ctx->peer.world;
        a.peer = peer_table(ctx, peer_slab);
        a.peer_dev = ctx->peer.tables.as<PeerTable>() + peer_slab;
        a.sim8 = nullptr; a.pitch8 = ctx->peer.pitch;
    }
    const uint64_t tiles = k12_num_tiles(ctx->n_own);
    CK(cudaMemsetAsync(ctx->scalars.p, 0, sizeof(unsigned long long) * SC_COUNT, ctx->stream), "memset scalars");
    if (emit) CK(cudaMemsetAsync(ctx->tile_state.p, 0, std::max<uint64_t>(tiles, 1) * 8, ctx->stream), "memset tile state");
    if (tiles) {
        LaunchScope ls(ctx, KID_K12_FUSED);
        CK(launch_k12(a, true, emit, exact32, ctx->stream), "k12_fused launch");
    }
    // the block path over the clusters set aside (it reads their count on the device), the carry
    // check, then ONE synchronisation for the whole pass
    if (tiles) {
        K2Args b;
        b.cl = nullptr; b.n_cl = 0; b.da = ctx->d_da; b.ebwt = a.ebwt;
        b.pos_base = ctx->pos_base; b.n_reads = ctx->n_reads; b.n_refs = ctx->n_refs;
        b.sim8 = a.sim8; b.pitch8 = a.pitch8; b.sim32 = a.sim32; b.lo = lo; b.hi = hi;
        b.peer = peer ? a.peer_dev : nullptr;
        b.carry = a.carry; b.sum_t = a.sum_t; b.large_count = a.long_count; b.large_list = nullptr; b.long_recs = a.long_list;
        b.large_cap = a.long_cap;
        b.large_next = scalar<unsigned long long>(ctx, SC_LARGE_NEXT);
        b.n_updates = a.n_updates;
        b.scratch = ctx->k2_scratch.as<uint8_t>(); b.scratch_per_cta = k2_block_scratch_bytes();
        {
            LaunchScope ls(ctx, KID_K2_BLOCK);
            CK(launch_k2_block(b, exact32, ctx->stream, k2_block_grid()), "k2_block launch");
        }
        if (a.sum_t && !peer && ctx->sim_rows * ctx->sim_pitch) {
            // The matrix was zero before this pass (lime_load) and every add is t << 8k on a 32-bit
            // word: a carry out of a byte lowers the byte sum by 255 (256 out of the top byte), so
            // byte sum == total of t  <=>  no cell exceeded 255.
            LaunchScope ls(ctx, KID_K12_CHECK);
            CK(launch_bytesum(ctx->d_sim, ctx->sim_rows * ctx->sim_pitch, scalar<unsigned long long>(ctx, SC_BYTESUM), ctx->stream),
               "k12_check launch");
        }
    }
    CK(cudaMemcpyAsync(sc, ctx->scalars.p, sizeof(unsigned long long) * SC_COUNT, cudaMemcpyDeviceToHost, ctx->stream), "D2H scalars");
    CK(cudaStreamSynchronize(ctx->stream), "k12_fused");
    if ((unsigned)sc[SC_ERR] & 1u)
        return fail(ctx, LIME_ERR_HALO, "lime_cluster_similarity: a cluster runs past the halo of this shard");
    if (sc[SC_MAXLEN] > LIME_SIZE_MAX_BUF)                       // ClusterBWT_DA.cpp:558-562
        return fail(ctx, LIME_ERR_MAXLEN, "maximum cluster size is greater than sizeMaxBuf");
    if (sc[SC_LARGE_COUNT] > a.long_cap)
        return fail(ctx, LIME_ERR_NOMEM, "lime_cluster_similarity: long-cluster list overflow");
    if (a.sum_t && !peer && sc[SC_SUMT] != sc[SC_BYTESUM]) sc[SC_CARRY] = 1;
    return LIME_OK;
}

extern "C" int lime_cluster_similarity(lime_ctx *ctx, uint32_t alpha, int ebwt_mode, int keep_clusters,
                                       uint64_t *n_clusters, uint64_t *max_len)
{
    if (!ctx) return LIME_ERR_ARG;
    if (!ctx->loaded) return fail(ctx, LIME_ERR_STATE, "lime_cluster_similarity: nothing loaded");
    if (!ctx->d_lcp && ctx->n_own) return fail(ctx, LIME_ERR_STATE, "lime_cluster_similarity: loaded without lcp");
    if (ctx->sim_done) return fail(ctx, LIME_ERR_STATE, "lime_cluster_similarity: already accumulated for this load");
    if (ebwt_mode && !ctx->d_ebwt) return fail(ctx, LIME_ERR_ARG, "lime_cluster_similarity: EBWT mode needs the ebwt array");
    CK(cudaSetDevice(ctx->device), "cudaSetDevice");
    const uint64_t n_total = ctx->n_own + ctx->n_halo;
    const uint64_t tiles = k12_num_tiles(ctx->n_own);
    for (int i = 0; i < 4; i++) ctx->sim_stats[i] = 0;
    ctx->have_clusters = false;
    CK(ctx->long_recs.reserve(k12_long_cap(n_total) * sizeof(lime_cluster)), "cudaMalloc long-cluster list");
    CK(ctx->k2_scratch.reserve(k2_block_scratch_bytes() * (uint64_t)k2_block_grid()), "cudaMalloc block-path scratch");
    const bool emit = keep_clusters != 0;
    if (emit) {
        CK(ctx->tile_state.reserve(std::max<uint64_t>(tiles, 1) * 8), "cudaMalloc tile state");
        const uint64_t cap = std::max<uint64_t>(ctx->cl_cap, std::max<uint64_t>(1024, n_total / 8));
        CK(ctx->clusters.reserve(cap * sizeof(lime_cluster)), "cudaMalloc clusters");
        ctx->cl_cap = ctx->clusters.cap / sizeof(lime_cluster);
    }
    unsigned long long sc[SC_COUNT];
    ctx->sim_done = true;
    int rc = run_k12_pass(ctx, alpha, ebwt_mode != 0, emit, false, 0, 0, sc, -1);
    if (rc) return rc;
    const uint64_t n_cl = sc[SC_NCLUSTERS], mx = sc[SC_MAXLEN];
    ctx->sim_stats[1] = sc[SC_LARGE_COUNT];
    ctx->sim_stats[0] = n_cl - sc[SC_LARGE_COUNT];
    ctx->sim_stats[2] = sc[SC_NUPDATES];
    ctx->n_clusters = n_cl; ctx->max_len = mx; ctx->alpha = alpha;
    if (n_clusters) *n_clusters = n_cl;
    if (max_len) *max_len = mx;
    if (emit) {
        if (n_cl > ctx->cl_cap) {
            // the records did not fit the first guess: produce them with a detection-only pass
            const bool was_done = ctx->sim_done;
            rc = lime_cluster_detect(ctx, alpha, nullptr, nullptr);
            ctx->sim_done = was_done;
            if (rc) return rc;
        }
        ctx->have_clusters = true;
    }
    if (((unsigned)sc[SC_CARRY]) == 0) return LIME_OK;

    // A packed add carried out of its byte: redo the pass into exact 32-bit cells, a band of reads
    // at a time, and store the sums mod 256 (uint8 wrap of SimArray_, ClusterBWT_DA.cpp:183,247).
    ctx->sim_stats[3] = 1;
    const uint64_t budget = 8ull << 30;
    uint64_t band = std::max<uint64_t>(1, budget / ((uint64_t)ctx->n_refs * 4));
    band = std::min<uint64_t>(band, ctx->n_reads);
    CK(ctx->sim32.reserve(band * ctx->n_refs * 4), "cudaMalloc exact accumulator");
    CK(cudaMemsetAsync(ctx->d_sim, 0, ctx->sim_rows * ctx->sim_pitch, ctx->stream), "memset similarity matrix");
    for (uint64_t lo = 0; lo < ctx->n_reads; lo += band) {
        const uint64_t hi = std::min<uint64_t>(lo + band, ctx->n_reads);
        CK(cudaMemsetAsync(ctx->sim32.p, 0, (hi - lo) * ctx->n_refs * 4, ctx->stream), "memset exact accumulator");
        rc = run_k12_pass(ctx, alpha, ebwt_mode != 0, false, true, (uint32_t)lo, (uint32_t)hi, sc, -1);
        if (rc) return rc;
        {
            LaunchScope ls(ctx, KID_K2_NARROW);
            CK(launch_k2_narrow(ctx->sim32.as<uint32_t>(), (uint32_t)lo, (uint32_t)hi, ctx->n_refs,
                                ctx->d_sim, ctx->sim_pitch, ctx->stream), "k2_narrow launch");
        }
    }
    CK(cudaStreamSynchronize(ctx->stream), "k12 exact pass");
    return LIME_OK;
}

extern "C" {

int lime_similarity(lime_ctx *ctx, int ebwt_mode)
{
    if (!ctx) return LIME_ERR_ARG;
    if (!ctx->loaded || !ctx->have_clusters) return fail(ctx, LIME_ERR_STATE, "lime_similarity: load and detect/set clusters first");
    if (ctx->sim_done) return fail(ctx, LIME_ERR_STATE, "lime_similarity: already accumulated for this load");
    if (ebwt_mode && !ctx->d_ebwt) return fail(ctx, LIME_ERR_ARG, "lime_similarity: EBWT mode needs the ebwt array");
    if (ctx->max_len > LIME_SIZE_MAX_BUF)                        // ClusterBWT_DA.cpp:558-562
        return fail(ctx, LIME_ERR_MAXLEN, "maximum cluster size is greater than sizeMaxBuf");
    CK(cudaSetDevice(ctx->device), "cudaSetDevice");
    for (int i = 0; i < 4; i++) ctx->sim_stats[i] = 0;
    ctx->sim_done = true;
    if (ctx->n_clusters == 0 || ctx->n_reads == 0 || ctx->n_refs == 0) return LIME_OK;
    CK(ctx->large_list.reserve(ctx->n_clusters * 8), "cudaMalloc block-path list");
    CK(ctx->k2_scratch.reserve(k2_block_scratch_bytes() * (uint64_t)k2_block_grid()), "cudaMalloc block-path scratch");

    int rc = run_k2_pass(ctx, ebwt_mode != 0, false, 0, 0);
    if (rc) return rc;
    unsigned long long sc[SC_COUNT];
    CK(cudaMemcpyAsync(sc, ctx->scalars.p, sizeof(sc), cudaMemcpyDeviceToHost, ctx->stream), "D2H scalars");
    CK(cudaStreamSynchronize(ctx->stream), "k2 similarity");
    ctx->sim_stats[1] = sc[SC_LARGE_COUNT];
    ctx->sim_stats[0] = ctx->n_clusters - sc[SC_LARGE_COUNT];
    ctx->sim_stats[2] = sc[SC_NUPDATES];
    if (((unsigned)sc[SC_CARRY]) == 0) return LIME_OK;

    // A packed add carried out of its byte: some cell's true sum exceeds 255.  Redo the pass into
    // exact 32-bit cells, a bounded band of reads at a time, and store the sums mod 256
    // (uint8 wrap of SimArray_, ClusterBWT_DA.cpp:183,247; Tools.h:68).
    ctx->sim_stats[3] = 1;
    const uint64_t budget = 8ull << 30;
    uint64_t band = std::max<uint64_t>(1, budget / ((uint64_t)ctx->n_refs * 4));
    band = std::min<uint64_t>(band, ctx->n_reads);
    CK(ctx->sim32.reserve(band * ctx->n_refs * 4), "cudaMalloc exact accumulator");
    // carries may also have spilled into the padding bytes of a row: start from a clean matrix
    CK(cudaMemsetAsync(ctx->d_sim, 0, ctx->sim_rows * ctx->sim_pitch, ctx->stream), "memset similarity matrix");
    for (uint64_t lo = 0; lo < ctx->n_reads; lo += band) {
        const uint64_t hi = std::min<uint64_t>(lo + band, ctx->n_reads);
        CK(cudaMemsetAsync(ctx->sim32.p, 0, (hi - lo) * ctx->n_refs * 4, ctx->stream), "memset exact accumulator");
        rc = run_k2_pass(ctx, ebwt_mode != 0, true, (uint32_t)lo, (uint32_t)hi);
        if (rc) return rc;
        {
            LaunchScope ls(ctx, KID_K2_NARROW);
            CK(launch_k2_narrow(ctx->sim32.as<uint32_t>(), (uint32_t)lo, (uint32_t)hi, ctx->n_refs,
                                ctx->d_sim, ctx->sim_pitch, ctx->stream), "k2_narrow launch");
        }
    }
    CK(cudaStreamSynchronize(ctx->stream), "k2 exact pass");
    return LIME_OK;
}

int lime_sim_stats(lime_ctx *ctx, uint64_t out[4])
{
    if (!ctx || !out) return LIME_ERR_ARG;
    for (int i = 0; i < 4; i++) out[i] = ctx->sim_stats[i];
    return LIME_OK;
}

int lime_sim_copy(lime_ctx *ctx, uint8_t *out)
{
    if (!ctx || !out) return LIME_ERR_ARG;
    if (!ctx->loaded) return fail(ctx, LIME_ERR_STATE, "lime_sim_copy: nothing loaded");
    if (ctx->n_reads == 0 || ctx->n_refs == 0) return LIME_OK;
    CK(cudaMemcpy2DAsync(out, ctx->n_refs, ctx->d_sim, ctx->sim_pitch, ctx->n_refs, ctx->n_reads,
                         cudaMemcpyDeviceToHost, ctx->stream), "D2H similarity matrix");
    CK(cudaStreamSynchronize(ctx->stream), "D2H similarity matrix");
    return LIME_OK;
}

int lime_choose(lime_ctx *ctx, int slot, uint32_t norm, float beta,
                uint32_t read_begin, uint32_t read_end, uint64_t *n_records)
{
    if (!ctx || slot < 0 || slot >= LIME_MAXFILES) return LIME_ERR_ARG;
    if (!ctx->loaded) return fail(ctx, LIME_ERR_STATE, "lime_choose: nothing loaded");
    if (read_begin > read_end || read_end > ctx->sim_rows) return fail(ctx, LIME_ERR_ARG, "lime_choose: bad read range");
    CK(cudaSetDevice(ctx->device), "cudaSetDevice");
    Slot &S = ctx->slots[slot];
    S.valid = false;
    const uint32_t n = read_end - read_begin;
    const uint64_t tiles = k3_num_tiles(n);
    CK(ctx->tile_state.reserve(std::max<uint64_t>(tiles, 1) * 8), "cudaMalloc tile state");
    CK(S.pos.reserve(std::max<uint64_t>(n, 1) * 8), "cudaMalloc pos");
    CK(cudaMemsetAsync(ctx->scalars.p, 0, sizeof(unsigned long long) * SC_COUNT, ctx->stream), "memset scalars");
    CK(cudaMemsetAsync(ctx->tile_state.p, 0, std::max<uint64_t>(tiles, 1) * 8, ctx->stream), "memset tile state");
    K3Args a;
    a.sim8 = ctx->d_sim; a.pitch8 = ctx->sim_pitch; a.n_refs = ctx->n_refs;
    a.read_begin = read_begin; a.read_end = read_end; a.norm = norm; a.beta = beta;
    a.tile_state = ctx->tile_state.as<unsigned long long>();
    a.tile_counter = scalar<unsigned int>(ctx, SC_TILE_COUNTER);
    a.n_records = scalar<unsigned long long>(ctx, SC_NRECORDS);
    a.pos = S.pos.as<uint64_t>(); a.bin = nullptr; a.bin_cap = 0;
    unsigned long long total = 1;
    // Every non-zero cell received at least one update, so 1 + reads + updates bounds the record
    // count and the single-pass kernel can run without a counting pass.  Not known for a matrix
    // bound by the caller: then count first, emit second.
    const bool single_pass = tiles && ctx->sim_done && !ctx->sim_bound && ctx->n_refs;
    if (single_pass) {
        const uint64_t cells = (uint64_t)n * ctx->n_refs;
        const uint64_t cap = 1 + (uint64_t)n + std::min<uint64_t>(ctx->sim_stats[2], cells);
        CK(S.bin.reserve(cap * sizeof(lime_pair_sim)), "cudaMalloc records");
        a.bin = S.bin.as<lime_pair_sim>(); a.bin_cap = S.bin.cap / sizeof(lime_pair_sim);
        {
            LaunchScope ls(ctx, KID_K3_CHOOSE);
            CK(launch_k3_choose(a, ctx->stream), "k3_choose launch");
        }
        CK(cudaMemcpyAsync(&total, a.n_records, 8, cudaMemcpyDeviceToHost, ctx->stream), "D2H record count");
        CK(cudaStreamSynchronize(ctx->stream), "k3_choose");
        if (total > a.bin_cap) return fail(ctx, LIME_ERR_STATE, "lime_choose: record bound violated");
    } else {
        if (tiles) {
            {
                LaunchScope ls(ctx, KID_K3_COUNT);
                CK(launch_k3_count(a, ctx->stream), "k3_count launch");
            }
            CK(cudaMemcpyAsync(&total, a.n_records, 8, cudaMemcpyDeviceToHost, ctx->stream), "D2H record count");
            CK(cudaStreamSynchronize(ctx->stream), "k3_count");
        }
        CK(S.bin.reserve(total * sizeof(lime_pair_sim)), "cudaMalloc records");
        a.bin = S.bin.as<lime_pair_sim>();
        {
            LaunchScope ls(ctx, KID_K3_EMIT);
            CK(launch_k3_emit(a, ctx->stream), "k3_emit launch");
        }
    }
    S.n_records = total; S.read_begin = read_begin; S.n_reads = n; S.valid = true;
    if (n_records) *n_records = total;
    return LIME_OK;
}

int lime_sim_export(lime_ctx *ctx, uint32_t read_begin, uint32_t read_end, void *dev_triples,
                    uint64_t cap, uint64_t *n_triples)
{
    if (!ctx || (!dev_triples && cap)) return LIME_ERR_ARG;
    if (!ctx->loaded) return fail(ctx, LIME_ERR_STATE, "lime_sim_export: nothing loaded");
    if (read_begin > read_end || read_end > ctx->sim_rows) return fail(ctx, LIME_ERR_ARG, "lime_sim_export: bad read range");
    CK(cudaSetDevice(ctx->device), "cudaSetDevice");
    const uint32_t n = read_end - read_begin;
    const uint64_t tiles = k3_num_tiles(n);
    unsigned long long total = 0;
    if (tiles) {
        CK(ctx->tile_state.reserve(tiles * 8), "cudaMalloc tile state");
        CK(cudaMemsetAsync(ctx->scalars.p, 0, sizeof(unsigned long long) * SC_COUNT, ctx->stream), "memset scalars");
        CK(cudaMemsetAsync(ctx->tile_state.p, 0, tiles * 8, ctx->stream), "memset tile state");
        K3Args a;
        a.sim8 = ctx->d_sim; a.pitch8 = ctx->sim_pitch; a.n_refs = ctx->n_refs;
        a.read_begin = read_begin; a.read_end = read_end; a.norm = 1; a.beta = 0.0f;
        a.tile_state = ctx->tile_state.as<unsigned long long>();
        a.tile_counter = scalar<unsigned int>(ctx, SC_TILE_COUNTER);
        a.n_records = scalar<unsigned long long>(ctx, SC_NRECORDS);
        a.pos = nullptr; a.bin = static_cast<lime_pair_sim *>(dev_triples); a.bin_cap = cap;
        {
            LaunchScope ls(ctx, KID_K3_EXPORT);
            CK(launch_k3_export(a, ctx->stream), "k3_export launch");
        }
        CK(cudaMemcpyAsync(&total, a.n_records, 8, cudaMemcpyDeviceToHost, ctx->stream), "D2H triple count");
        CK(cudaStreamSynchronize(ctx->stream), "k3_export");
    }
    if (n_triples) *n_triples = total;
    if (total > cap) return fail(ctx, LIME_ERR_NOMEM, "lime_sim_export: triple buffer too small");
    return LIME_OK;
}

int lime_sim_import(lime_ctx *ctx, const void *dev_triples, uint64_t n)
{
    if (!ctx || (!dev_triples && n)) return LIME_ERR_ARG;
    if (!ctx->loaded) return fail(ctx, LIME_ERR_STATE, "lime_sim_import: nothing loaded");
    CK(cudaSetDevice(ctx->device), "cudaSetDevice");
    if (n == 0) return LIME_OK;
    CK(cudaMemsetAsync(scalar<unsigned int>(ctx, SC_ERR), 0, 4, ctx->stream), "memset err");
    {
        LaunchScope ls(ctx, KID_K3_IMPORT);
        CK(launch_k3_import(static_cast<const lime_pair_sim *>(dev_triples), n, ctx->d_sim, ctx->sim_pitch,
                            (uint32_t)std::min<uint64_t>(ctx->sim_rows, 0xffffffffull), ctx->n_refs,
                            scalar<unsigned int>(ctx, SC_ERR), ctx->stream), "k3_import launch");
    }
    unsigned err = 0;
    CK(cudaMemcpyAsync(&err, scalar<unsigned int>(ctx, SC_ERR), 4, cudaMemcpyDeviceToHost, ctx->stream), "D2H err");
    CK(cudaStreamSynchronize(ctx->stream), "k3_import");
    if (err & 1u) return fail(ctx, LIME_ERR_ARG, "lime_sim_import: triple outside the matrix");
    ctx->sim_stats[2] += n;            // keeps the record bound of lime_choose valid
    return LIME_OK;
}

int lime_choose_copy(lime_ctx *ctx, int slot, lime_pair_sim *bin, uint64_t *pos)
{
    if (!ctx || slot < 0 || slot >= LIME_MAXFILES) return LIME_ERR_ARG;
    Slot &S = ctx->slots[slot];
    if (!S.valid) return fail(ctx, LIME_ERR_STATE, "lime_choose_copy: empty slot");
    if (bin) CK(cudaMemcpyAsync(bin, S.bin.p, S.n_records * sizeof(lime_pair_sim), cudaMemcpyDeviceToHost, ctx->stream), "D2H records");
    if (pos && S.n_reads) CK(cudaMemcpyAsync(pos, S.pos.p, (uint64_t)S.n_reads * 8, cudaMemcpyDeviceToHost, ctx->stream), "D2H pos");
    CK(cudaStreamSynchronize(ctx->stream), "D2H slot");
    return LIME_OK;
}

int lime_slot_set(lime_ctx *ctx, int slot, const lime_pair_sim *bin, uint64_t n_records,
                  const uint64_t *pos, uint32_t n_reads)
{
    if (!ctx || slot < 0 || slot >= LIME_MAXFILES || (!bin && n_records) || (!pos && n_reads)) return LIME_ERR_ARG;
    CK(cudaSetDevice(ctx->device), "cudaSetDevice");
    Slot &S = ctx->slots[slot];
    S.valid = false;
    CK(S.bin.reserve(std::max<uint64_t>(n_records, 1) * sizeof(lime_pair_sim)), "cudaMalloc records");
    CK(S.pos.reserve(std::max<uint64_t>(n_reads, 1) * 8), "cudaMalloc pos");
    if (n_records) CK(cudaMemcpyAsync(S.bin.p, bin, n_records * sizeof(lime_pair_sim), cudaMemcpyHostToDevice, ctx->stream), "H2D records");
    if (n_reads) CK(cudaMemcpyAsync(S.pos.p, pos, (uint64_t)n_reads * 8, cudaMemcpyHostToDevice, ctx->stream), "H2D pos");
    CK(cudaStreamSynchronize(ctx->stream), "H2D slot");
    S.n_records = n_records; S.read_begin = 0; S.n_reads = n_reads; S.valid = true;
    return LIME_OK;
}

} // extern "C"
