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

int reserve_score_scratch(lime_ctx *ctx)
{
    const uint64_t n_total = ctx->n_own + ctx->n_halo;
    CK(ctx->long_recs.reserve(k12_long_cap(n_total) * sizeof(lime_cluster)), "cudaMalloc long-cluster list");
    CK(ctx->k2_scratch.reserve(k2_block_scratch_bytes() * (uint64_t)k2_block_grid()), "cudaMalloc block-path scratch");
    CK(ctx->strand_state.reserve(sizeof(unsigned long long) * SS_COUNT * (LIME_MAXFILES + 1)), "cudaMalloc strand state");
    if (!ctx->strand_host)
        CK(cudaHostAlloc((void **)&ctx->strand_host, sizeof(unsigned long long) * SS_COUNT * (LIME_MAXFILES + 1), cudaHostAllocDefault),
           "cudaHostAlloc strand state");
    return LIME_OK;
}

// One fused pass (steps 1 + 2a) plus the block path over the clusters it set aside (it reads their count on
// the device).  Nothing synchronises here; the scalars land in the device block `sc` (layout SS_*), which
// the caller has zeroed.  flat.rec != null: the updates leave as records (multi-GPU) instead of being added
// to ctx->d_sim.
int queue_k12_score(lime_ctx *ctx, uint32_t alpha, bool ebwt, unsigned long long *sc, const FlatOut &flat, int max_ctas_per_sm)
{
    { const int rc = check_alpha(ctx, alpha); if (rc) return rc; }
    K12Args a = {};
    a.lcp = ctx->d_lcp; a.lcp8 = ctx->lcp8 ? 1 : 0; a.da = ctx->d_da; a.ebwt = ebwt ? ctx->d_ebwt : nullptr;
    a.n_own = ctx->n_own; a.n_total = ctx->n_own + ctx->n_halo; a.pos_base = ctx->pos_base; a.is_last = ctx->is_last;
    a.alpha = alpha; a.n_reads = ctx->n_reads; a.n_refs = ctx->n_refs;
    a.n_clusters = sc + SS_NCLUSTERS; a.max_len = sc + SS_MAXLEN;
    a.err = reinterpret_cast<unsigned int *>(sc + SS_ERR);
    a.long_list = ctx->long_recs.as<lime_cluster>();
    a.long_count = sc + SS_LONG_COUNT;
    a.long_cap = ctx->long_recs.cap / sizeof(lime_cluster);
    a.sim8 = ctx->d_sim; a.pitch8 = ctx->sim_pitch;
    a.n_updates = sc + SS_NUPDATES; a.sum_t = sc + SS_SUMT;
    a.flat = flat.rec; a.flat_count = flat.count; a.flat_cap = flat.cap;
    a.pol_evict_first = ctx->l2_policy[0]; a.pol_evict_last = ctx->l2_policy[1];
    a.max_ctas_per_sm = max_ctas_per_sm;
    if (k12_num_tiles(ctx->n_own) == 0) return LIME_OK;
    {
        LaunchScope ls(ctx, KID_K12_FUSED);
        CK(launch_k12(a, flat.rec ? K12_FLAT : K12_DIRECT, ctx->stream), "k12_fused launch");
    }
    K2Args b = {};
    b.cl = nullptr; b.n_cl = 0; b.da = ctx->d_da; b.ebwt = a.ebwt;
    b.pos_base = ctx->pos_base; b.n_reads = ctx->n_reads; b.n_refs = ctx->n_refs;
    b.sim8 = a.sim8; b.pitch8 = a.pitch8; b.sim32 = nullptr; b.lo = 0; b.hi = 0;
    // the block path inspects the word every add returns: a carry there is reported like one of the fused
    // kernel, by a byte sum that falls short of sum_t (it adds its t to sum_t only when the add did not carry)
    b.carry = reinterpret_cast<unsigned int *>(sc + SS_ERR) + 1;
    b.sum_t = a.sum_t; b.large_count = a.long_count; b.large_list = nullptr; b.long_recs = a.long_list;
    b.large_cap = a.long_cap;
    b.large_next = sc + SS_LONG_NEXT;
    b.n_updates = a.n_updates;
    b.scratch = ctx->k2_scratch.as<uint8_t>(); b.scratch_per_cta = k2_block_scratch_bytes();
    b.flat = flat;
    {
        LaunchScope ls(ctx, KID_K2_BLOCK);
        CK(launch_k2_block(b, false, ctx->stream, k2_block_grid()), "k2_block launch");
    }
    return LIME_OK;
}

// exact redo of the accumulation after a carry: cluster list + 32-bit cells for a band of reads at a time,
// stored mod 256 (uint8 wrap of SimArray_, ClusterBWT_DA.cpp:183,247; Tools.h:68)
static int exact_bands(lime_ctx *ctx, bool ebwt)
{
    ctx->sim_stats[3] = 1;
    const uint64_t budget = 8ull << 30;
    uint64_t band = std::max<uint64_t>(1, budget / std::max<uint64_t>((uint64_t)ctx->n_refs * 4, 1));
    band = std::min<uint64_t>(band, ctx->n_reads);
    CK(ctx->sim32.reserve(std::max<uint64_t>(band * ctx->n_refs * 4, 16)), "cudaMalloc exact accumulator");
    CK(ctx->large_list.reserve(std::max<uint64_t>(ctx->n_clusters, 1) * 8), "cudaMalloc block-path list");
    CK(ctx->k2_scratch.reserve(k2_block_scratch_bytes() * (uint64_t)k2_block_grid()), "cudaMalloc block-path scratch");
    // carries may also have spilled into the padding bytes of a row: start from a clean matrix
    CK(cudaMemsetAsync(ctx->d_sim, 0, ctx->sim_rows * ctx->sim_pitch, ctx->stream), "memset similarity matrix");
    for (uint64_t lo = 0; lo < ctx->n_reads; lo += band) {
        const uint64_t hi = std::min<uint64_t>(lo + band, ctx->n_reads);
        CK(cudaMemsetAsync(ctx->sim32.p, 0, (hi - lo) * ctx->n_refs * 4, ctx->stream), "memset exact accumulator");
        const int rc = run_k2_pass(ctx, ebwt, true, (uint32_t)lo, (uint32_t)hi);
        if (rc) return rc;
        {
            LaunchScope ls(ctx, KID_K2_NARROW);
            CK(launch_k2_narrow(ctx->sim32.as<uint32_t>(), (uint32_t)lo, (uint32_t)hi, ctx->n_refs,
                                ctx->d_sim, ctx->sim_pitch, ctx->stream), "k2_narrow launch");
        }
    }
    CK(cudaStreamSynchronize(ctx->stream), "exact pass");
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
    if (ctx->xchg.active) return fail(ctx, LIME_ERR_STATE, "lime_cluster_similarity: the context is in exchange mode (lime_xchg_accumulate)");
    CK(cudaSetDevice(ctx->device), "cudaSetDevice");
    for (int i = 0; i < 4; i++) ctx->sim_stats[i] = 0;
    ctx->have_clusters = false;
    ctx->sim_clean = false;
    if ((uint64_t)ctx->n_reads + ctx->n_refs >= K12_ID_LIMIT) {
        // more documents than the fused kernel's packed stream element holds: cluster list + list kernels
        int rc = lime_cluster_detect(ctx, alpha, n_clusters, max_len);
        if (rc) return rc;
        return lime_similarity(ctx, ebwt_mode);
    }
    int rc = reserve_score_scratch(ctx);
    if (rc) return rc;
    unsigned long long *sc = strand_scalar<unsigned long long>(ctx, LIME_MAXFILES, 0);
    unsigned long long *hs = ctx->strand_host + (size_t)LIME_MAXFILES * SS_COUNT;
    CK(cudaMemsetAsync(sc, 0, sizeof(unsigned long long) * SS_COUNT, ctx->stream), "memset strand state");
    ctx->sim_done = true;
    rc = queue_k12_score(ctx, alpha, ebwt_mode != 0, sc, FlatOut());
    if (rc) return rc;
    if (ctx->sim_rows * ctx->sim_pitch && k12_num_tiles(ctx->n_own)) {
        // The matrix was zero before this pass (lime_load) and every add is t << 8k on a 32-bit word: a carry
        // out of a byte lowers the byte sum by 255 (256 out of the top byte), so
        // byte sum == total of t  <=>  no cell exceeded 255.
        LaunchScope ls(ctx, KID_K12_CHECK);
        CK(launch_bytesum(ctx->d_sim, ctx->sim_rows * ctx->sim_pitch, sc + SS_BYTESUM, ctx->stream), "k12_check launch");
    }
    CK(cudaMemcpyAsync(hs, sc, sizeof(unsigned long long) * SS_COUNT, cudaMemcpyDeviceToHost, ctx->stream), "D2H strand state");
    CK(cudaStreamSynchronize(ctx->stream), "k12_fused");
    if (hs[SS_ERR] & 1ull)
        return fail(ctx, LIME_ERR_HALO, "lime_cluster_similarity: a cluster runs past the halo of this shard");
    if (hs[SS_MAXLEN] > LIME_SIZE_MAX_BUF)                       // ClusterBWT_DA.cpp:558-562
        return fail(ctx, LIME_ERR_MAXLEN, "maximum cluster size is greater than sizeMaxBuf");
    const uint64_t n_cl = hs[SS_NCLUSTERS], mx = hs[SS_MAXLEN];
    ctx->sim_stats[1] = hs[SS_LONG_COUNT];
    ctx->sim_stats[0] = n_cl - hs[SS_LONG_COUNT];
    ctx->sim_stats[2] = hs[SS_NUPDATES];
    ctx->n_clusters = n_cl; ctx->max_len = mx; ctx->alpha = alpha;
    if (n_clusters) *n_clusters = n_cl;
    if (max_len) *max_len = mx;
    const bool carry = hs[SS_SUMT] != hs[SS_BYTESUM] || (hs[SS_ERR] >> 32) != 0;
    if (keep_clusters || carry) {
        // the records (.clrs) come from the detection passes (count, scan, emit); same chunk ownership
        unsigned long long dsc[SC_COUNT];
        rc = detect_into_ctx(ctx, alpha, dsc);
        if (rc) return rc;
        if (dsc[SC_NCLUSTERS] != n_cl) return fail(ctx, LIME_ERR_STATE, "lime_cluster_similarity: detection and scoring disagree");
        ctx->have_clusters = true;
    }
    if (!carry) return LIME_OK;
    // A cell exceeded 255: redo the accumulation exactly from the cluster list.
    rc = exact_bands(ctx, ebwt_mode != 0);
    if (rc) return rc;
    if (!keep_clusters) ctx->have_clusters = false;
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
    ctx->sim_clean = false;
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

    // A packed add carried out of its byte: some cell's true sum exceeds 255.
    return exact_bands(ctx, ebwt_mode != 0);
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
    if (!ctx->d_sim) return fail(ctx, LIME_ERR_STATE, "lime_sim_copy: no matrix selected");
    // exchange mode: the selected slab, i.e. the rows of the reads this rank owns
    const uint64_t rows = ctx->xchg.active ? ctx->sim_rows : ctx->n_reads;
    { const int rc = xchg_join(ctx); if (rc) return rc; }
    CK(cudaMemcpy2DAsync(out, ctx->n_refs, ctx->d_sim, ctx->sim_pitch, ctx->n_refs, rows,
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
    S.valid = false; S.deferred = false;
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
    S.bin_cap = S.bin.cap / sizeof(lime_pair_sim);
    S.status[0] = ctx->n_clusters; S.status[1] = ctx->max_len; S.status[2] = total; S.status[3] = ctx->sim_stats[2];
    ctx->rec_hint = std::max<uint64_t>(ctx->rec_hint, total);
    if (n_records) *n_records = total;
    return LIME_OK;
}

} // extern "C"

// clusterChoose of rows [0, n) of ctx->d_sim into `slot` without a host synchronisation: the single-pass kernel
// with a record buffer of `cap` records, the byte sum of the rows (carry proof) and the record count left in
// the strand's device state; the rows are left zero (the next strand finds a clean matrix).
int queue_choose_deferred(lime_ctx *ctx, int slot, uint32_t norm, float beta, uint32_t n, uint64_t cap, int kind)
{
    Slot &S = ctx->slots[slot];
    S.valid = false;
    const uint64_t tiles = k3_num_tiles(n);
    CK(ctx->tile_state.reserve(std::max<uint64_t>(tiles, 1) * 8), "cudaMalloc tile state");
    CK(S.pos.reserve(std::max<uint64_t>(n, 1) * 8), "cudaMalloc pos");
    CK(S.bin.reserve(std::max<uint64_t>(cap, 1) * sizeof(lime_pair_sim)), "cudaMalloc records");
    CK(cudaMemsetAsync(ctx->tile_state.p, 0, std::max<uint64_t>(tiles, 1) * 8, ctx->stream), "memset tile state");
    K3Args a;
    a.sim8 = ctx->d_sim; a.pitch8 = ctx->sim_pitch; a.n_refs = ctx->n_refs;
    a.read_begin = 0; a.read_end = n; a.norm = norm; a.beta = beta;
    a.tile_state = ctx->tile_state.as<unsigned long long>();
    a.tile_counter = strand_scalar<unsigned int>(ctx, slot, SS_TILE_COUNTER);
    a.n_records = strand_scalar<unsigned long long>(ctx, slot, SS_NRECORDS);
    a.pos = S.pos.as<uint64_t>(); a.bin = S.bin.as<lime_pair_sim>(); a.bin_cap = S.bin.cap / sizeof(lime_pair_sim);
    a.byte_sum = strand_scalar<unsigned long long>(ctx, slot, SS_BYTESUM);
    a.zero_after = 1;
    if (tiles) {
        LaunchScope ls(ctx, KID_K3_CHOOSE);
        CK(launch_k3_choose(a, ctx->stream), "k3_choose launch");
    } else {
        // no reads: the record list is the sentinel alone
        const unsigned long long one = 1;
        CK(cudaMemcpyAsync(a.n_records, &one, 8, cudaMemcpyHostToDevice, ctx->stream), "H2D record count");
        CK(cudaMemsetAsync(S.bin.p, 0, sizeof(lime_pair_sim), ctx->stream), "memset sentinel");
    }
    S.n_records = 0; S.read_begin = 0; S.n_reads = n; S.valid = true; S.deferred = true;
    S.bin_cap = a.bin_cap; S.kind = kind;
    ctx->sim_clean = true; ctx->sim_clean_ptr = ctx->d_sim; ctx->sim_clean_bytes = ctx->sim_rows * ctx->sim_pitch;
    return LIME_OK;
}

extern "C" {

int lime_strand_run(lime_ctx *ctx, int slot, uint32_t alpha, int ebwt_mode, uint32_t norm, float beta)
{
    if (!ctx || slot < 0 || slot >= LIME_MAXFILES) return LIME_ERR_ARG;
    if (!ctx->loaded) return fail(ctx, LIME_ERR_STATE, "lime_strand_run: nothing loaded");
    if (ctx->xchg.active) return fail(ctx, LIME_ERR_STATE, "lime_strand_run: the context is in exchange mode");
    const bool list_path = (uint64_t)ctx->n_reads + ctx->n_refs >= K12_ID_LIMIT;
    if (ctx->rec_hint == 0 || ctx->sim_bound || list_path || ctx->n_reads == 0 || ctx->n_refs == 0) {
        // first strand of this context (no size for the record buffer yet) and the corner cases: the
        // synchronising calls
        uint64_t n_cl = 0, mx = 0, n_rec = 0;
        int rc = lime_cluster_similarity(ctx, alpha, ebwt_mode, 0, &n_cl, &mx);
        if (rc) return rc;
        rc = lime_choose(ctx, slot, norm, beta, 0, ctx->n_reads, &n_rec);
        if (rc) return rc;
        ctx->slots[slot].status[0] = n_cl; ctx->slots[slot].status[1] = mx;
        return LIME_OK;
    }
    if (!ctx->d_lcp && ctx->n_own) return fail(ctx, LIME_ERR_STATE, "lime_strand_run: loaded without lcp");
    if (ctx->sim_done) return fail(ctx, LIME_ERR_STATE, "lime_strand_run: already accumulated for this load");
    if (ebwt_mode && !ctx->d_ebwt) return fail(ctx, LIME_ERR_ARG, "lime_strand_run: EBWT mode needs the ebwt array");
    CK(cudaSetDevice(ctx->device), "cudaSetDevice");
    int rc = reserve_score_scratch(ctx);
    if (rc) return rc;
    for (int i = 0; i < 4; i++) ctx->sim_stats[i] = 0;
    ctx->have_clusters = false;
    unsigned long long *sc = strand_scalar<unsigned long long>(ctx, slot, 0);
    CK(cudaMemsetAsync(sc, 0, sizeof(unsigned long long) * SS_COUNT, ctx->stream), "memset strand state");
    ctx->sim_done = true;
    ctx->sim_clean = false;
    rc = queue_k12_score(ctx, alpha, ebwt_mode != 0, sc, FlatOut());
    if (rc) return rc;
    const uint64_t cells = (uint64_t)ctx->n_reads * ctx->n_refs;
    const uint64_t cap = 1 + (uint64_t)ctx->n_reads + std::min<uint64_t>(cells, ctx->rec_hint + ctx->rec_hint / 4 + 4096);
    return queue_choose_deferred(ctx, slot, norm, beta, ctx->n_reads, cap, 1);
}

int lime_strand_status(lime_ctx *ctx, int slot, uint64_t out[4])
{
    if (!ctx || !out || slot < 0 || slot >= LIME_MAXFILES) return LIME_ERR_ARG;
    const Slot &S = ctx->slots[slot];
    if (!S.valid) return fail(ctx, LIME_ERR_STATE, "lime_strand_status: empty slot");
    if (S.deferred) return fail(ctx, LIME_ERR_STATE, "lime_strand_status: the strand has not been read back yet (lime_classify does)");
    for (int i = 0; i < 4; i++) out[i] = S.status[i];
    return LIME_OK;
}

int lime_sim_export(lime_ctx *ctx, uint32_t read_begin, uint32_t read_end, void *dev_triples,
                    uint64_t cap, uint64_t *n_triples)
{
    if (!ctx || (!dev_triples && cap)) return LIME_ERR_ARG;
    if (!ctx->loaded) return fail(ctx, LIME_ERR_STATE, "lime_sim_export: nothing loaded");
    if (read_begin > read_end || read_end > ctx->sim_rows) return fail(ctx, LIME_ERR_ARG, "lime_sim_export: bad read range");
    if (ctx->n_refs > (1u << 24)) return fail(ctx, LIME_ERR_ARG, "lime_sim_export: a triple holds genome << 8 | count, n_refs must not exceed 2^24");
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
    if (ctx->n_refs > (1u << 24)) return fail(ctx, LIME_ERR_ARG, "lime_sim_import: a triple holds genome << 8 | count, n_refs must not exceed 2^24");
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
    ctx->sim_clean = false;
    return LIME_OK;
}

int lime_choose_copy(lime_ctx *ctx, int slot, lime_pair_sim *bin, uint64_t *pos)
{
    if (!ctx || slot < 0 || slot >= LIME_MAXFILES) return LIME_ERR_ARG;
    Slot &S = ctx->slots[slot];
    if (!S.valid) return fail(ctx, LIME_ERR_STATE, "lime_choose_copy: empty slot");
    { const int rc = xchg_join(ctx); if (rc) return rc; }
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
