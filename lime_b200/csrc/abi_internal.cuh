// abi_internal.cuh -- helpers shared by the ABI translation units.
#pragma once
#include "common.cuh"
#include <string>

using namespace lime;


static const char *const kKernelNames[KID_COUNT] = {
    "k12_fused", "k12_check", "k1_detect", "k2_warp", "k2_block", "k2_narrow", "k3_choose", "k3_export", "k3_import", "k3_count", "k3_emit", "k3_classify", "x_route", "x_apply", "x_publish", "x_wait"};

// device scalar block layout (u64 slots)
enum {
    SC_NCLUSTERS = 0, SC_MAXLEN, SC_LARGE_COUNT, SC_LARGE_NEXT, SC_NUPDATES, SC_NRECORDS,
    SC_TILE_COUNTER,   // u32 used
    SC_ERR,            // u32 used
    SC_CARRY,          // u32 used
    SC_SUMT, SC_BYTESUM,
    SC_COUNT
};

// per-strand device state of the calls that do not synchronise (lime_strand_run, lime_xchg_*): one block of
// SS_COUNT u64 per strand slot in ctx->strand_state; read back once, by lime_classify / lime_xchg_status
enum {
    SS_NCLUSTERS = 0, SS_MAXLEN, SS_LONG_COUNT, SS_LONG_NEXT, SS_NUPDATES, SS_SUMT, SS_BYTESUM, SS_NRECORDS,
    SS_ERR,            // bit0 halo exhausted, bit2 inbox region overflow, bit3 record out of range
    SS_TILE_COUNTER,   // u32 used
    SS_FLAT_COUNT, SS_APPLIED_T, SS_APPLIED_N,
    SS_SENT = 16,      // [XCHG_MAX][XCHG_SUB_MAX] records routed to every (owner, row block)
    SS_COUNT = 160
};
static_assert(SS_SENT + lime::XCHG_MAX * lime::XCHG_SUB_MAX <= SS_COUNT, "strand state block");

static int fail(lime_ctx *c, int code, const std::string &msg)
{
    if (c) c->err = msg;
    return code;
}
static int fail_cuda(lime_ctx *c, cudaError_t e, const char *what)
{
    return fail(c, e == cudaErrorMemoryAllocation ? LIME_ERR_NOMEM : LIME_ERR_CUDA,
                std::string(what) + ": " + cudaGetErrorString(e));
}
#define CK(call, what)                                                     \
    do { cudaError_t e__ = (call); if (e__ != cudaSuccess) return fail_cuda(ctx, e__, what); } while (0)

template <class T> static T *scalar(lime_ctx *c, int slot)
{
    return reinterpret_cast<T *>(c->scalars.as<unsigned long long>() + slot);
}

template <class T> static T *strand_scalar(lime_ctx *c, int slot, int k)
{
    return reinterpret_cast<T *>(c->strand_state.as<unsigned long long>() + (size_t)slot * SS_COUNT + k);
}

// lime_abi_steps.cu / lime_abi_xchg.cu
// the fused kernel + the block path over the clusters it sets aside, asynchronous; scalars land in `sc`
// (a device block laid out as SS_*); flat: records instead of adds into ctx->d_sim
int queue_k12_score(lime_ctx *ctx, uint32_t alpha, bool ebwt, unsigned long long *sc, const lime::FlatOut &flat, int max_ctas_per_sm = 0);
int reserve_score_scratch(lime_ctx *ctx);
int queue_choose_deferred(lime_ctx *ctx, int slot, uint32_t norm, float beta, uint32_t n, uint64_t cap, int kind);
int detect_into_ctx(lime_ctx *ctx, uint32_t alpha, unsigned long long host_sc[SC_COUNT]);
void xchg_release(lime_ctx *ctx);
// the context's stream waits for everything queued on the exchange's side stream (no-op without an exchange)
int xchg_join(lime_ctx *ctx);
int strand_readback(lime_ctx *ctx, int slot);
// lime_abi_load.cu
int load_host_narrow(lime_ctx *ctx, const uint32_t *lcp, const uint32_t *da, const uint8_t *ebwt, uint64_t n, uint32_t bits_alpha, bool da3);
// LIME_ERR_STATE when the resident lcp cannot answer lcp < alpha (narrowed to bytes and alpha > 255, or packed for another alpha)
int check_alpha(lime_ctx *ctx, uint32_t alpha);
void load_release(lime_ctx *ctx);

// bracket one kernel launch with events when profiling
struct LaunchScope {
    lime_ctx *c; int kid; cudaEvent_t a = nullptr, b = nullptr;
    LaunchScope(lime_ctx *c_, int kid_) : c(c_), kid(kid_)
    {
        c->launches++;
        c->prof_launches[kid]++;
        if (c->prof_on) {
            cudaEventCreate(&a); cudaEventCreate(&b);
            cudaEventRecord(a, c->stream);
        }
    }
    ~LaunchScope()
    {
        if (a) { cudaEventRecord(b, c->stream); c->prof_events.push_back({kid, a, b}); }
    }
};

