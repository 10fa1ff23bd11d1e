// abi_internal.cuh -- helpers shared by the ABI translation units.
#pragma once
#include "common.cuh"
#include <string>

using namespace lime;


static const char *const kKernelNames[KID_COUNT] = {
    "k12_fused", "k12_check", "k1_detect", "k2_warp", "k2_block", "k2_narrow", "k3_choose", "k3_export", "k3_import", "k3_count", "k3_emit", "k3_classify", "load_widen"};

// device scalar block layout (u64 slots)
enum {
    SC_NCLUSTERS = 0, SC_MAXLEN, SC_LARGE_COUNT, SC_LARGE_NEXT, SC_NUPDATES, SC_NRECORDS,
    SC_TILE_COUNTER,   // u32 used
    SC_ERR,            // u32 used
    SC_CARRY,          // u32 used
    SC_SUMT, SC_BYTESUM,
    SC_COUNT
};

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

// lime_abi_steps.cu / lime_abi_peer.cu
int run_k12_pass(lime_ctx *ctx, uint32_t alpha, bool ebwt, bool emit, bool exact32, uint32_t lo, uint32_t hi,
                 unsigned long long sc[SC_COUNT], int peer_slab);
lime::PeerTable peer_table(const lime_ctx *ctx, int slab);
void peer_release(lime_ctx *ctx);
// lime_abi_load.cu
int load_host_narrow(lime_ctx *ctx, const uint32_t *lcp, const uint32_t *da, const uint8_t *ebwt, uint64_t n);
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

