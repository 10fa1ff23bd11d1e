// lime_abi_classify.cu -- ABI entry points for step 3 and the instrumentation getters.
#include "abi_internal.cuh"
#include <algorithm>
#include <vector>

// Evaluate the device state of a strand that was run without a synchronisation (ctx->strand_host holds the copy).
// LIME_ERR_RETRY: the result of that strand is not usable (record buffer too small, or a similarity cell
// exceeded 255 and wrapped where the reference wraps, ClusterBWT_DA.cpp:183,247): redo it with the
// synchronising calls, which size exactly and take the exact path.
int strand_readback(lime_ctx *ctx, int slot)
{
    Slot &S = ctx->slots[slot];
    const unsigned long long *h = ctx->strand_host + (size_t)slot * SS_COUNT;
    S.deferred = false;
    S.n_records = h[SS_NRECORDS];
    S.status[0] = h[SS_NCLUSTERS]; S.status[1] = h[SS_MAXLEN]; S.status[2] = h[SS_NRECORDS]; S.status[3] = h[SS_NUPDATES];
    ctx->rec_hint = std::max<uint64_t>(ctx->rec_hint, h[SS_NRECORDS]);
    if (h[SS_ERR] & 1ull) { S.valid = false; return fail(ctx, LIME_ERR_HALO, "a cluster runs past the halo of this shard"); }
    if (h[SS_MAXLEN] > LIME_SIZE_MAX_BUF) { S.valid = false; return fail(ctx, LIME_ERR_MAXLEN, "maximum cluster size is greater than sizeMaxBuf"); }
    if (h[SS_ERR] & 16ull) { S.valid = false; return fail(ctx, LIME_ERR_STATE, "exchange: a source rank did not deliver its records within 20 s"); }
    if (h[SS_ERR] & 12ull) { S.valid = false; return fail(ctx, (h[SS_ERR] & 8ull) ? LIME_ERR_ARG : LIME_ERR_RETRY, (h[SS_ERR] & 8ull) ? "exchange: update record out of range" : "exchange: an inbox region overflowed"); }
    if (S.kind == 2 && h[SS_FLAT_COUNT] > ctx->xchg.flat_cap) { S.valid = false; return fail(ctx, LIME_ERR_RETRY, "exchange: the update list overflowed"); }
    const unsigned long long added = S.kind == 2 ? h[SS_APPLIED_T] : h[SS_SUMT];
    if (added != h[SS_BYTESUM] || (h[SS_ERR] >> 32) != 0) { S.valid = false; return fail(ctx, LIME_ERR_RETRY, "deferred strand: a similarity cell exceeded 255"); }
    if (h[SS_NRECORDS] > S.bin_cap) { S.valid = false; return fail(ctx, LIME_ERR_RETRY, "deferred strand: record buffer too small"); }
    return LIME_OK;
}

extern "C" {

int lime_classify(lime_ctx *ctx, int n_files, uint32_t n_refs,
                  const uint32_t *v_corRef, const uint32_t *m_corRef, int tax_rank, int higher,
                  uint8_t *out_type, uint32_t *out_tax, float *out_sim)
{
    if (!ctx) return LIME_ERR_ARG;
    if (n_files != 2 && n_files != 4)                            // Classify.cpp:334-338
        return fail(ctx, LIME_ERR_ARG, "the allowed number of input files is 2 (single-end reads), or 4 (paired-end reads)");
    if (!v_corRef && n_refs) return fail(ctx, LIME_ERR_ARG, "lime_classify: v_corRef is NULL");
    if (higher && tax_rank < 1) return fail(ctx, LIME_ERR_ARG, "lime_classify: HIGHER needs tax_rank >= 1");
    if (tax_rank < 0 || tax_rank > LIME_RANKS) return fail(ctx, LIME_ERR_ARG, "lime_classify: tax_rank out of range");
    for (int i = 0; i < n_files; i++) {
        if (!ctx->slots[i].valid) return fail(ctx, LIME_ERR_STATE, "lime_classify: strand slot not filled");
        if (ctx->slots[i].n_reads != ctx->slots[0].n_reads || ctx->slots[i].read_begin != ctx->slots[0].read_begin)
            return fail(ctx, LIME_ERR_STATE, "lime_classify: strand slots cover different read ranges");
    }
    CK(cudaSetDevice(ctx->device), "cudaSetDevice");
    { const int rc = xchg_join(ctx); if (rc) return rc; }        // strands finished on the exchange's side stream
    const uint32_t n_reads = ctx->slots[0].n_reads;
    const bool use_m = higher && m_corRef;

    // taxonomy tables -> device
    const uint64_t tab_words = (uint64_t)n_refs * (1 + (use_m ? LIME_RANKS : 0));
    CK(ctx->tax_tables.reserve(std::max<uint64_t>(tab_words, 1) * 4), "cudaMalloc taxonomy");
    uint32_t *d_v = ctx->tax_tables.as<uint32_t>();
    uint32_t *d_m = use_m ? d_v + n_refs : nullptr;
    if (n_refs) CK(cudaMemcpyAsync(d_v, v_corRef, (uint64_t)n_refs * 4, cudaMemcpyHostToDevice, ctx->stream), "H2D v_corRef");
    if (use_m && n_refs) CK(cudaMemcpyAsync(d_m, m_corRef, (uint64_t)n_refs * 4 * LIME_RANKS, cudaMemcpyHostToDevice, ctx->stream), "H2D m_corRef");

    ClsArgs a;
    a.n_files = n_files;
    for (int i = 0; i < LIME_MAXFILES; i++) {
        const bool on = i < n_files;
        a.bin[i] = on ? ctx->slots[i].bin.as<lime_pair_sim>() : nullptr;
        a.pos[i] = on ? ctx->slots[i].pos.as<uint64_t>() : nullptr;
        a.n_records[i] = on ? ctx->slots[i].n_records : 0;
        const bool dev = on && ctx->slots[i].deferred;
        a.n_records_dev[i] = dev ? strand_scalar<unsigned long long>(ctx, i, SS_NRECORDS) : nullptr;
        a.bin_cap[i] = on ? ctx->slots[i].bin_cap : 0;
    }
    a.n_reads = n_reads; a.n_refs = n_refs; a.v = d_v; a.m = d_m;
    a.tax_rank = tax_rank; a.higher = use_m ? 1 : 0;
    // summaries over the whole database for the stage-C corner where every genome qualifies
    a.all_v_same = 1; a.all_v_tax = n_refs ? v_corRef[0] : 0;
    for (uint32_t g = 1; g < n_refs; g++) if (v_corRef[g] != v_corRef[0]) { a.all_v_same = 0; break; }
    for (int x = 0; x < LIME_RANKS; x++) {
        a.all_m_same[x] = 1; a.all_m_tax[x] = 0;
        if (use_m && n_refs) {
            const uint32_t *row = m_corRef + (size_t)x * n_refs;
            a.all_m_tax[x] = row[0];
            for (uint32_t g = 1; g < n_refs; g++) if (row[g] != row[0]) { a.all_m_same[x] = 0; break; }
        }
    }
    // outputs: type u8[n] | tax u32[n] | sim f32[n]; keep 16-byte sections
    const uint64_t n16 = ((uint64_t)n_reads + 15) / 16 * 16;
    CK(ctx->cls_out.reserve(std::max<uint64_t>(n16 * 9, 16)), "cudaMalloc classify outputs");
    uint8_t *base = ctx->cls_out.as<uint8_t>();
    a.out_tax = reinterpret_cast<uint32_t *>(base);
    a.out_sim = reinterpret_cast<float *>(base + n16 * 4);
    a.out_type = base + n16 * 8;
    a.err = scalar<unsigned int>(ctx, SC_ERR);
    CK(cudaMemsetAsync(ctx->scalars.p, 0, sizeof(unsigned long long) * SC_COUNT, ctx->stream), "memset scalars");
    if (n_reads) {
        LaunchScope ls(ctx, KID_K3_CLASSIFY);
        CK(launch_classify(a, ctx->stream), "k3_classify launch");
    }
    unsigned err = 0;
    CK(cudaMemcpyAsync(&err, a.err, 4, cudaMemcpyDeviceToHost, ctx->stream), "D2H err");
    bool any_deferred = false;
    for (int i = 0; i < n_files; i++) any_deferred |= ctx->slots[i].deferred;
    if (any_deferred)
        CK(cudaMemcpyAsync(ctx->strand_host, ctx->strand_state.p, sizeof(unsigned long long) * SS_COUNT * LIME_MAXFILES,
                           cudaMemcpyDeviceToHost, ctx->stream), "D2H strand state");
    if (n_reads) {
        if (out_tax) CK(cudaMemcpyAsync(out_tax, a.out_tax, (uint64_t)n_reads * 4, cudaMemcpyDeviceToHost, ctx->stream), "D2H tax");
        if (out_sim) CK(cudaMemcpyAsync(out_sim, a.out_sim, (uint64_t)n_reads * 4, cudaMemcpyDeviceToHost, ctx->stream), "D2H sim");
        if (out_type) CK(cudaMemcpyAsync(out_type, a.out_type, n_reads, cudaMemcpyDeviceToHost, ctx->stream), "D2H type");
    }
    CK(cudaStreamSynchronize(ctx->stream), "k3_classify");
    // the strands that ran without a host synchronisation: their scalars and error flags, read back just now
    int deferred_rc = LIME_OK;
    for (int i = 0; i < n_files; i++) {
        Slot &S = ctx->slots[i];
        if (!S.deferred) continue;
        const int rc = strand_readback(ctx, i);
        if (rc && !deferred_rc) deferred_rc = rc;
    }
    if (deferred_rc) { ctx->cls_n_valid = false; return deferred_rc; }
    ctx->cls_n = n_reads; ctx->cls_n_valid = !(err & 1u);
    if (err & 1u) return fail(ctx, LIME_ERR_ROWS, "lime_classify: a .res row is not strictly ascending in genome id (or out of range)");
    return LIME_OK;
}

int lime_classify_result(lime_ctx *ctx, uint8_t *type, uint32_t *tax, float *sim, int mem)
{
    if (!ctx) return LIME_ERR_ARG;
    if (!ctx->cls_n_valid) return fail(ctx, LIME_ERR_STATE, "lime_classify_result: no classification to return");
    CK(cudaSetDevice(ctx->device), "cudaSetDevice");
    const uint64_t n = ctx->cls_n, n16 = (n + 15) / 16 * 16;
    const uint8_t *base = ctx->cls_out.as<uint8_t>();
    const cudaMemcpyKind kind = mem == LIME_MEM_DEVICE ? cudaMemcpyDeviceToDevice : cudaMemcpyDeviceToHost;
    if (n) {
        if (tax) CK(cudaMemcpyAsync(tax, base, n * 4, kind, ctx->stream), "copy tax");
        if (sim) CK(cudaMemcpyAsync(sim, base + n16 * 4, n * 4, kind, ctx->stream), "copy sim");
        if (type) CK(cudaMemcpyAsync(type, base + n16 * 8, n, kind, ctx->stream), "copy type");
    }
    if (mem != LIME_MEM_DEVICE) CK(cudaStreamSynchronize(ctx->stream), "lime_classify_result");
    return LIME_OK;
}

uint64_t lime_launch_count(const lime_ctx *ctx) { return ctx ? ctx->launches : 0; }

int lime_prof_enable(lime_ctx *ctx, int on)
{
    if (!ctx) return LIME_ERR_ARG;
    ctx->prof_on = on != 0;
    return LIME_OK;
}

int lime_prof_kernels(void) { return KID_COUNT; }

static int prof_drain(lime_ctx *ctx)
{
    if (ctx->prof_events.empty()) return LIME_OK;
    CK(cudaStreamSynchronize(ctx->stream), "prof sync");
    for (auto &p : ctx->prof_events) {
        float ms = 0.f;
        if (cudaEventElapsedTime(&ms, p.a, p.b) == cudaSuccess) ctx->prof_ms[p.kid] += ms;
        cudaEventDestroy(p.a); cudaEventDestroy(p.b);
    }
    ctx->prof_events.clear();
    return LIME_OK;
}

int lime_prof_get(lime_ctx *ctx, int k, const char **name, uint64_t *launches, double *ms)
{
    if (!ctx || k < 0 || k >= KID_COUNT) return LIME_ERR_ARG;
    int rc = prof_drain(ctx);
    if (rc) return rc;
    if (name) *name = kKernelNames[k];
    if (launches) *launches = ctx->prof_launches[k];
    if (ms) *ms = ctx->prof_ms[k];
    return LIME_OK;
}

int lime_prof_reset(lime_ctx *ctx)
{
    if (!ctx) return LIME_ERR_ARG;
    int rc = prof_drain(ctx);
    for (int k = 0; k < KID_COUNT; k++) { ctx->prof_launches[k] = 0; ctx->prof_ms[k] = 0; }
    return rc;
}

} // extern "C"
