// common.cuh -- shared declarations of the lime-b200 CUDA translation units (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <string>
#include <vector>
#include "../../include/lime_types.h"
#include "../../include/lime_cuda.h"

namespace lime {

// ---------------------------------------------------------------------------------------------
// small device helpers
// ---------------------------------------------------------------------------------------------
constexpr unsigned FULL = 0xffffffffu;

__device__ __forceinline__ unsigned lane_id() { return threadIdx.x & 31u; }

// 128-bit read-only streaming load (inputs are read once per kernel)
__device__ __forceinline__ uint4 ldg128(const uint32_t *p)
{
    return __ldg(reinterpret_cast<const uint4 *>(p));
}

// 128-bit streaming load, evict-first in L1/L2: keeps the similarity accumulator's lines in L2
__device__ __forceinline__ uint4 ldcs128(const uint32_t *p)
{
    return __ldcs(reinterpret_cast<const uint4 *>(p));
}

__device__ __forceinline__ unsigned long long ld_relaxed_u64(const unsigned long long *p)
{
    unsigned long long v;
    asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_relaxed_u64(unsigned long long *p, unsigned long long v)
{
    asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}

// ---- decoupled look-back over 64-bit tile descriptors ----------------------------------------
// descriptor = status (top 2 bits) | value (low 62 bits); the array is zeroed (INVALID) before
// the launch and tiles are numbered by an atomic ticket so every predecessor is already running.
constexpr unsigned long long LB_AGG = 1ull << 62;      // tile aggregate published
constexpr unsigned long long LB_PREFIX = 2ull << 62;   // inclusive prefix published
constexpr unsigned long long LB_VALUE = (1ull << 62) - 1;

// Called by all 32 lanes of one warp.  Returns the exclusive prefix of `agg` over tiles < tile.
__device__ __forceinline__ unsigned long long lookback_exclusive(unsigned long long *state,
                                                                 unsigned tile,
                                                                 unsigned long long agg)
{
    const unsigned lane = lane_id();
    if (tile == 0) {
        if (lane == 0) st_relaxed_u64(&state[0], LB_PREFIX | agg);
        return 0;
    }
    if (lane == 0) st_relaxed_u64(&state[tile], LB_AGG | agg);
    unsigned long long excl = 0;
    long long idx = (long long)tile - 1;
    while (true) {
        long long t = idx - lane;
        unsigned long long d = LB_PREFIX;   // virtual tile before 0: prefix 0
        if (t >= 0) {
            d = ld_relaxed_u64(&state[t]);
            while ((d >> 62) == 0) d = ld_relaxed_u64(&state[t]);
        }
        unsigned pm = __ballot_sync(FULL, (d >> 62) == 2);
        int first = pm ? (__ffs(pm) - 1) : 32;
        unsigned long long v = ((int)lane <= first) ? (d & LB_VALUE) : 0ull;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(FULL, v, o);
        excl += v;
        if (pm) break;
        idx -= 32;
    }
    if (lane == 0) st_relaxed_u64(&state[tile], LB_PREFIX | ((excl + agg) & LB_VALUE));
    return excl;
}

// ---------------------------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------------------------
constexpr int LIME_STAGE_SLOTS = 3;
constexpr uint64_t LIME_STAGE_CHUNK = 8ull << 20;   // positions per chunk of the narrowed lcp transfer

struct DevBuf {
    void *p = nullptr;
    size_t cap = 0;
    cudaError_t reserve(size_t bytes)
    {
        if (bytes <= cap) return cudaSuccess;
        if (p) cudaFree(p);
        p = nullptr; cap = 0;
        cudaError_t e = cudaMalloc(&p, bytes);
        if (e == cudaSuccess) cap = bytes;
        return e;
    }
    void release() { if (p) cudaFree(p); p = nullptr; cap = 0; }
    template <class T> T *as() const { return reinterpret_cast<T *>(p); }
};

enum KernelId {
    KID_K12_FUSED = 0,
    KID_K12_CHECK,
    KID_K1_DETECT,
    KID_K2_WARP,
    KID_K2_BLOCK,
    KID_K2_NARROW,
    KID_K3_CHOOSE,
    KID_K3_EXPORT,
    KID_K3_IMPORT,
    KID_K3_COUNT,
    KID_K3_EMIT,
    KID_K3_CLASSIFY,
    KID_X_ROUTE,
    KID_X_APPLY,
    KID_X_PUBLISH,
    KID_X_WAIT,
    KID_COUNT
};

struct Slot {
    DevBuf bin, pos;
    uint64_t n_records = 0;
    uint32_t read_begin = 0, n_reads = 0;
    bool valid = false;
    // filled by a call that did not synchronise: n_records lives in the strand's device state until
    // lime_classify has read it back; status = nClusters, maxLen, records, updates of that strand
    bool deferred = false;
    uint64_t bin_cap = 0;
    uint64_t status[4] = {0, 0, 0, 0};
    int kind = 0;                   // deferred: 1 = one GPU (carry proof: sum_t), 2 = exchange (applied_t)
};

struct ProfEvent { int kid; cudaEvent_t a, b; };

// multi-GPU exchange of similarity updates (lime_abi_xchg.cu): the rows of a read live on its owner only
constexpr int XCHG_MAX = 16;
constexpr int XCHG_SUB_MAX = 8;            // row blocks an owner's slab is applied in (each small enough for L2)
constexpr int XCHG_MAIL_STRIDE = XCHG_SUB_MAX + 1;   // uint64 per (strand, source) in an inbox header: counts + flag
struct XchgState {
    bool active = false, loopback = false;
    int world = 0, rank = 0, n_strands = 0;
    uint32_t per = 0;                      // reads per owner
    uint64_t rows = 0, pitch = 0;          // per slab
    uint64_t flat_cap = 0, inbox_cap = 0;  // records; inbox_cap per (source, strand) region = n_sub * sub_cap
    uint32_t n_sub = 1, sub_rows = 1;      // row blocks of a slab and rows per block
    uint64_t sub_cap = 0;
    DevBuf slabs;                          // [n_strands][rows][pitch] uint8: the similarity rows of the owned reads
    DevBuf flat[2];                        // [flat_cap] records {read, genome << 8 | t}; strand s uses list s & 1, so
                                           // that the next strand is scored while this one's list is being routed
    // inbox: a mailbox header (mail_bytes: [n_strands][world][XCHG_MAIL_STRIDE] uint64 = the record counts of
    // source p per row block + the epoch flag p raises after its records have landed), then the regions
    // [n_strands][world][inbox_cap] records; plain cudaMalloc (CUDA IPC)
    void *inbox_own = nullptr;
    void *inbox_map[XCHG_MAX] = {nullptr}; // every rank's inbox in this process (own included)
    uint64_t mail_bytes = 0;
    uint64_t epoch[LIME_MAXFILES] = {0};   // accumulations of strand s so far = the flag value its sources raise
    // routing, applying and clusterChoose of strand s are queued on `tail` and run beside the scoring of strand
    // s + 1 on the context's stream (the scoring kernel leaves k12_spare CTA slots of every SM to them)
    cudaStream_t tail = nullptr;
    cudaEvent_t ev_scored[LIME_MAXFILES] = {nullptr};   // main: list of strand s complete
    cudaEvent_t ev_routed[2] = {nullptr};                // tail: list s & 1 has been read
    cudaEvent_t ev_strand[LIME_MAXFILES] = {nullptr};    // tail: last operation queued for strand s
    cudaEvent_t ev_join = nullptr;
    bool routed_set[2] = {false, false}, strand_set[LIME_MAXFILES] = {false};
    bool tail_dirty = false, use_tail = true;
    int k12_ctas = 0;                      // CTAs per SM of the scoring kernel while tail work may run (0 = all)
    DevBuf state;                          // device scalars, see lime_abi_xchg.cu
};

} // namespace lime

struct lime_ctx {
    int device = 0;
    cudaStream_t own_stream = nullptr, stream = nullptr;
    std::string err;

    // resident strand
    const void *d_lcp = nullptr;            // uint32 values, or -- lcp8 -- the bytes min(lcp, 255) of a narrowed load
    bool lcp8 = false;
    const uint32_t *d_da = nullptr;
    const uint8_t *d_ebwt = nullptr;
    lime::DevBuf in_lcp, in_da, in_ebwt;
    uint64_t n_own = 0, n_halo = 0, pos_base = 0;
    int is_last = 1;
    uint32_t n_reads = 0, n_refs = 0;
    bool loaded = false;
    // narrowed lcp transfer (lime_abi_load.cu): alpha_limit is the caller's promise (lime_set_alpha_limit);
    // such a load leaves min(lcp, 255) as bytes on the device (lcp8) and the kernels read those
    uint32_t alpha_limit = 0;
    lime::DevBuf in_lcp8;
    // lime_set_alpha: the caller names the alpha itself; a host load then ships the bit lcp < alpha and leaves the
    // bytes 0 / 255 in in_lcp8; lcp_alpha = the alpha the resident lcp answers for (0: any alpha <= 255, or u32)
    uint32_t alpha_exact = 0, lcp_alpha = 0;
    lime::DevBuf in_lcpbits, in_da3;
    uint64_t stage_bytes = 0;               // size of each page-locked stage buffer
    bool da3_loaded = false;                // the last host load shipped da as 3 bytes per position
    uint8_t *stage[lime::LIME_STAGE_SLOTS] = {nullptr};
    cudaEvent_t stage_ev[lime::LIME_STAGE_SLOTS] = {nullptr};
    bool stage_busy = false;

    // clusters
    lime::DevBuf clusters;          // lime_cluster[cl_cap]
    uint64_t cl_cap = 0, n_clusters = 0, max_len = 0;
    uint32_t alpha = 0;
    bool have_clusters = false;

    // step-1 scratch
    lime::DevBuf tile_state;        // u64 per tile
    lime::DevBuf scalars;           // small block of device scalars (see abi)
    uint64_t l2_policy[2] = {0, 0}; // createpolicy words: evict_first, evict_last (k12_policies, lime_init)
    lime::DevBuf strand_state;      // [LIME_MAXFILES][SS_COUNT] u64: scalars of the strands run without a host sync
    unsigned long long *strand_host = nullptr;   // page-locked mirror of strand_state
    uint64_t rec_hint = 0;          // records of the largest strand result seen (sizes the next deferred one)
    bool sim_clean = false;         // the matrix is known to be all zero (left so by the choose pass)
    const void *sim_clean_ptr = nullptr; uint64_t sim_clean_bytes = 0;

    // similarity accumulator
    lime::DevBuf sim_own;
    uint8_t *d_sim = nullptr;       // [sim_rows][sim_pitch] uint8
    uint64_t sim_rows = 0, sim_pitch = 0;
    bool sim_bound = false, sim_done = false;
    lime::DevBuf sim32;             // exact pass scratch
    lime::DevBuf large_list;        // u64 cluster indices for the block path
    lime::DevBuf long_recs;         // fused path: records of the clusters left to the block path
    lime::DevBuf k2_scratch;        // block-path hash tables
    uint64_t sim_stats[4] = {0, 0, 0, 0};

    // step-2/3 results
    lime::Slot slots[LIME_MAXFILES];
    lime::DevBuf row_counts;
    lime::DevBuf cls_out;           // classify outputs
    uint64_t cls_n = 0; bool cls_n_valid = false;
    lime::DevBuf tax_tables;        // v_corRef + m_corRef on device
    lime::XchgState xchg;

    // instrumentation
    uint64_t launches = 0;
    bool prof_on = false;
    std::vector<lime::ProfEvent> prof_events;
    uint64_t prof_launches[lime::KID_COUNT] = {0};
    double prof_ms[lime::KID_COUNT] = {0};
};

namespace lime {

// launchers implemented in the kernel translation units; all asynchronous on ctx->stream

// records {read, genome << 8 | t} of similarity updates, appended instead of added when the rows of the
// matrix live on other GPUs (multi-GPU exchange, lime_abi_xchg.cu / xchg.cu)
struct FlatOut {
    uint2 *rec = nullptr;                 // null: updates go into the matrix
    unsigned long long *count = nullptr;  // slots handed out so far (every slot below it holds a record, t = 0: none)
    uint64_t cap = 0;                     // slots the list can hold; more -> the host reports LIME_ERR_NOMEM
};

struct K2Args {
    const lime_cluster *cl;
    uint64_t n_cl;
    const uint32_t *da;
    const uint8_t *ebwt;              // null in DA-only mode
    uint64_t pos_base;
    uint32_t n_reads, n_refs;
    // packed uint8 accumulator
    uint8_t *sim8; uint64_t pitch8;
    // exact 32-bit accumulator for reads [lo, hi)
    uint32_t *sim32; uint32_t lo, hi;
    unsigned int *carry;              // set to 1 when a packed add carried out of its byte
    unsigned long long *sum_t;        // when non-null: total of all t added (see K12Args::sum_t)
    unsigned long long *large_count;  // block-path work list
    unsigned long long *large_list;   // indices into cl (cluster-list path) ...
    const lime_cluster *long_recs;    // ... or, when non-null, the records themselves (fused path)
    uint64_t large_cap;               // entries the work list can hold
    unsigned long long *large_next;   // ticket for the block path
    unsigned long long *n_updates;    // statistics
    uint8_t *scratch; uint64_t scratch_per_cta;
    FlatOut flat;                     // block path only
};
cudaError_t launch_k2_warp(const K2Args &a, bool exact32, cudaStream_t s);
cudaError_t launch_k2_block(const K2Args &a, bool exact32, cudaStream_t s, int grid);
uint64_t k2_block_scratch_bytes();
int k2_block_grid();
cudaError_t launch_k2_narrow(const uint32_t *sim32, uint32_t lo, uint32_t hi, uint32_t n_refs,
                             uint8_t *sim8, uint64_t pitch8, cudaStream_t s);

// fused steps 1 + 2a (k12_stream.cu)
constexpr uint32_t K12_ID_LIMIT = 0x07ffffffu;   // the stream kernel packs id | class << 27 | start << 31
constexpr int K12_FLAT_BLK = 256;                // slots a warp takes from the flat list at a time
enum { K12_COUNT = 0, K12_DETECT = 1, K12_DIRECT = 2, K12_FLAT = 3 };
struct K12Args {
    const void *lcp;                  // uint32[n_total], or uint8[n_total] = min(lcp, 255) when lcp8
    int lcp8;
    const uint32_t *da;
    const uint8_t *ebwt;              // null in DA-only mode
    uint64_t n_own, n_total, pos_base;
    int is_last;
    uint32_t alpha, n_reads, n_refs;
    lime_cluster *out;                // K12_DETECT: records in ascending pStart order
    uint64_t out_cap;
    unsigned long long *tile_state;   // K12_DETECT: per-chunk cluster counts, turned into record offsets by a scan
    unsigned long long *n_clusters;   // device scalars
    unsigned long long *max_len;
    unsigned int *err;                // bit0: halo exhausted
    lime_cluster *long_list;          // clusters the window path does not take
    unsigned long long *long_count;
    uint64_t long_cap;
    uint8_t *sim8; uint64_t pitch8;   // K12_DIRECT: packed uint8 accumulator, updated with fire-and-forget adds
    unsigned long long *n_updates;
    // total of all t issued, so that the absence of byte carries can be proven (byte sum of the matrix == sum_t)
    unsigned long long *sum_t;
    uint2 *flat;                      // K12_FLAT: the record list (FlatOut)
    unsigned long long *flat_count;
    uint64_t flat_cap;
    // L2 policy words (createpolicy): inputs evict_first, matrix lines evict_last; from k12_policies()
    uint64_t pol_evict_first, pol_evict_last;
    int max_ctas_per_sm;              // > 0: leave the other CTA slots of an SM to kernels of another stream
};
cudaError_t k12_policies(uint64_t out[2]);
cudaError_t launch_bytesum(const uint8_t *p, uint64_t bytes, unsigned long long *out, cudaStream_t s);
cudaError_t launch_k12(const K12Args &a, int what, cudaStream_t s);
uint64_t k12_num_tiles(uint64_t n_own);
uint64_t k12s_num_chunks(uint64_t n_own);
uint64_t k12_long_cap(uint64_t n_total);

// multi-GPU routing of the update records (xchg.cu)
struct XchgRoute {
    const uint2 *flat; const unsigned long long *flat_count; uint64_t flat_cap;
    uint32_t world, per, n_refs;
    uint32_t n_sub, sub_rows;
    uint2 *dst[XCHG_MAX];             // owner o's inbox region for this source and strand (peer or local memory)
    uint64_t sub_cap;
    unsigned long long *sent;         // [world][n_sub] device counters (local)
    unsigned long long *err;
};
// after route_kernel: this source's counts for owner o -> mail[o][0 .. n_sub), then the flag mail[o][XCHG_SUB_MAX] = epoch
struct XchgPublish {
    const unsigned long long *sent;   // [world][n_sub]
    unsigned long long *mail[XCHG_MAX];
    uint32_t world, n_sub;
    unsigned long long epoch;
};
// on the owner: until every source has raised `epoch` in mail[p][XCHG_SUB_MAX] (bounded: err bit 4 after 20 s)
struct XchgWait {
    const unsigned long long *mail;   // this strand's [world][XCHG_MAIL_STRIDE]
    uint32_t world; int only;         // only >= 0: that source alone (loopback)
    unsigned long long epoch;
    unsigned long long *err;
};
struct XchgApply {
    const uint2 *src; uint64_t sub_cap;
    const unsigned long long *mail;   // this strand's [world][XCHG_MAIL_STRIDE]: records source p sent for row block j
    int only;                         // >= 0: apply that source's region alone (loopback)
    uint32_t world, rank, n_sub, sub;
    uint8_t *slab; uint64_t pitch; uint32_t rows, n_refs;
    unsigned long long *applied_t, *applied_n, *err;
};
cudaError_t launch_xchg_route(const XchgRoute &a, cudaStream_t s);
cudaError_t launch_xchg_apply(const XchgApply &a, cudaStream_t s);
cudaError_t launch_xchg_publish(const XchgPublish &a, cudaStream_t s);
cudaError_t launch_xchg_wait(const XchgWait &a, cudaStream_t s);

struct K3Args {
    uint8_t *sim8; uint64_t pitch8;
    uint32_t n_refs, read_begin, read_end;
    uint32_t norm; float beta;
    unsigned long long *tile_state;
    unsigned int *tile_counter;
    unsigned long long *n_records;    // device scalar: records incl. sentinel
    uint64_t *pos;                    // [read_end-read_begin]
    lime_pair_sim *bin;               // emit only
    uint64_t bin_cap;                 // single-pass kernel: capacity of bin (writes are guarded)
    // single-pass kernel, optional: add the byte sum of the rows to *byte_sum (carry proof of the fire-and-forget
    // adds, one pass over the matrix instead of a separate one) and leave the rows zero for the next strand
    unsigned long long *byte_sum = nullptr;
    int zero_after = 0;
};
cudaError_t launch_k3_choose(const K3Args &a, cudaStream_t s);
cudaError_t launch_k3_export(const K3Args &a, cudaStream_t s);
cudaError_t launch_k3_import(const lime_pair_sim *tr, uint64_t n, uint8_t *sim8, uint64_t pitch8, uint32_t n_reads,
                             uint32_t n_refs, unsigned int *err, cudaStream_t s);
cudaError_t launch_k3_count(const K3Args &a, cudaStream_t s);
cudaError_t launch_k3_emit(const K3Args &a, cudaStream_t s);
uint64_t k3_num_tiles(uint32_t n_reads);

struct ClsArgs {
    int n_files;
    const lime_pair_sim *bin[LIME_MAXFILES];
    const uint64_t *pos[LIME_MAXFILES];
    uint64_t n_records[LIME_MAXFILES];
    // a strand that was run without a host synchronisation: its record count is still on the device; bin_cap
    // bounds what may be read there
    const unsigned long long *n_records_dev[LIME_MAXFILES];
    uint64_t bin_cap[LIME_MAXFILES];
    uint32_t n_reads, n_refs;
    const uint32_t *v;                // [n_refs]
    const uint32_t *m;                // [6][n_refs] or null
    int tax_rank, higher;
    // all-genome summaries for the "every genome is near the maximum" corner of Exam_2
    uint32_t all_v_same, all_v_tax;
    uint32_t all_m_same[LIME_RANKS], all_m_tax[LIME_RANKS];
    uint8_t *out_type; uint32_t *out_tax; float *out_sim;
    unsigned int *err;                // bit0: row not strictly ascending / out of range
};
cudaError_t launch_classify(const ClsArgs &a, cudaStream_t s);

} // namespace lime
