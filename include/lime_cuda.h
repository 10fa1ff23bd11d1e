/*
 * lime_cuda.h -- the C ABI of lime-b200: LiME's classification hot path on one B200.
 *
 * The reference (veronicaguerrini/LiME) has no plugin / FFI interface: the path is three
 * executables that talk through files, and inside each executable one OpenMP parallel
 * region does the work.  This ABI puts a seam exactly at those regions, so a maintainer
 * can replace the body of each region with one call (INTEGRATION.md shows the patch):
 *
 *   lime_cluster_detect   <->  ClusterLCP.cpp:139-283     (#pragma omp parallel scan of lcp/da)
 *   lime_similarity       <->  ClusterBWT_DA.cpp:630-670  (clusterAnalyze, :256-358)
 *   lime_choose           <->  ClusterBWT_DA.cpp:737      (clusterChoose, :361-450)
 *   lime_classify         <->  Classify.cpp:474-702       (#pragma omp parallel for over reads)
 *
 * Conventions
 *   - extern "C", plain pointers and sizes; type widths are those of Tools.h (lime_types.h).
 *   - every function returns 0 on success or a negative LIME_ERR_* code; the message is
 *     available from lime_last_error(ctx).  Nothing throws across the boundary.
 *   - one lime_ctx per process and GPU (one process per GPU); a ctx is not thread-safe.
 *   - pointers are HOST pointers unless the parameter is documented as device memory or the
 *     call takes a `mem` argument (LIME_MEM_HOST / LIME_MEM_DEVICE).
 *   - there is no CPU fallback: without a CUDA device lime_init fails with LIME_ERR_CUDA.
 */
#ifndef LIME_CUDA_H_
#define LIME_CUDA_H_

#include <stdint.h>
#include "lime_types.h"

#ifdef __cplusplus
extern "C" {
#endif

#define LIME_OK            0
#define LIME_ERR_ARG      -1   /* bad argument / precondition                                   */
#define LIME_ERR_CUDA     -2   /* CUDA runtime error (message has the cudaError string)          */
#define LIME_ERR_MAXLEN   -3   /* a cluster is longer than sizeMaxBuf (ClusterBWT_DA.cpp:558-562) */
#define LIME_ERR_HALO     -4   /* a straddling cluster runs past the supplied halo               */
#define LIME_ERR_STATE    -5   /* call order violated (e.g. similarity before cluster_detect)    */
#define LIME_ERR_NOMEM    -6   /* device or host allocation failed                               */
#define LIME_ERR_ROWS     -7   /* .res rows not strictly ascending in genome id                  */
#define LIME_ERR_RETRY    -8   /* a strand run without synchronisation must be redone with the   */
                               /* synchronising calls (buffer guess too small, or a cell > 255)  */

#define LIME_MEM_HOST      0
#define LIME_MEM_DEVICE    1

typedef struct lime_ctx lime_ctx;

/* ---- context ------------------------------------------------------------------------------ */

/* Create a context on CUDA device `device` (cudaSetDevice ordinal). */
int  lime_init(lime_ctx **ctx, int device);
void lime_destroy(lime_ctx *ctx);
const char *lime_last_error(const lime_ctx *ctx);      /* ctx may be NULL: last init error */
const char *lime_version(void);

/* Launch all work of this ctx on an existing cudaStream_t (passed as void*), e.g. the caller's
 * current stream so that its CUDA events bracket our kernels.  NULL restores the ctx's own stream;
 * pass cudaStreamLegacy ((void*)0x1) for the default stream. */
int  lime_set_stream(lime_ctx *ctx, void *cuda_stream);

/* Page-locked host memory for buffers handed to lime_load / lime_classify / the copy functions:
 * transfers from / to it run at full PCIe rate and asynchronously.  NULL on failure. */
void *lime_pinned_alloc(uint64_t bytes);
void  lime_pinned_free(void *p);

/* ---- inputs ------------------------------------------------------------------------------- */

/*
 * Make one strand's arrays resident.  lcp/da have n_own + n_halo elements, ebwt likewise or NULL
 * (EBWT=0 build, ClusterBWT_DA.cpp:16-18).  Element 0 is global position pos_base.
 *   n_own    positions this GPU owns: a cluster belongs to the shard that holds the position
 *            with lcp < alpha which opens it -- the thread-chunk rule of ClusterLCP.cpp:150-202;
 *   n_halo   positions after the owned range that a straddling cluster may run into (:246-264);
 *   is_last  non-zero if element n_own+n_halo-1 is the last position of the whole array (:244-245).
 * Single GPU: n_halo = 0, pos_base = 0, is_last = 1.
 * mem = LIME_MEM_HOST copies (pageable or pinned) host arrays to the device.  From page-locked memory the
 * copies are asynchronous: lime_load may return while they are in flight, so the arrays must not be modified
 * or freed before the next synchronising call on this ctx (lime_cluster_detect, lime_cluster_similarity,
 * lime_similarity, lime_choose, lime_classify, any *_copy) has returned;
 * mem = LIME_MEM_DEVICE borrows 16-byte aligned device pointers, which must stay valid.
 * Loading resets the clusters and zeroes the similarity accumulator.
 */
int  lime_load(lime_ctx *ctx, const uint32_t *lcp, const uint32_t *da, const uint8_t *ebwt,
               uint64_t n_own, uint64_t n_halo, uint64_t pos_base, int is_last,
               uint32_t n_reads, uint32_t n_refs, int mem);

/*
 * Narrowed lcp transfer.  Every step reads lcp only through `lcp[i] < alpha` (ClusterLCP.cpp:150-202,
 * :246-264), so for alpha <= 255 the byte min(lcp, 255) decides every test as the 32-bit value does.
 * lime_set_alpha_limit declares the largest alpha the caller will use on the strands it loads from now
 * on (the drop-in executables know it from their command line before they read a file); 0 = no promise
 * (default).  With 1 <= max_alpha <= 255, lime_load(LIME_MEM_HOST) narrows lcp on host threads, chunk by
 * chunk into page-locked stage buffers while the previous chunks are on the bus, sends 1 + 4 (+1) instead
 * of 4 + 4 (+1) bytes per position over PCIe and the kernels read the bytes; results are identical.  A step
 * called with alpha > 255 on a strand loaded that way fails with LIME_ERR_STATE.
 */
int  lime_set_alpha_limit(lime_ctx *ctx, uint32_t max_alpha);
/*
 * One step further: lime_set_alpha names the alpha itself (the drop-in executables have it on their command
 * line).  With 1 <= alpha <= 255, lime_load(LIME_MEM_HOST) ships ONE BIT per position -- lcp[i] < alpha, packed by
 * the same host threads, bit j of byte i = position 8 i + j -- i.e. 1/8 + 4 (+1) bytes per position, and expands
 * each chunk on the device into the bytes 0 (lcp < alpha) / 255 the kernels read.  Results are identical for that
 * alpha; a step called with another alpha on a strand loaded that way fails with LIME_ERR_STATE.  0 = off (default);
 * it takes precedence over lime_set_alpha_limit.  In this mode da crosses as 3 bytes per position while every document
 * id fits (n_reads + n_refs <= 2^24; packed by the same host threads, expanded on the device into the uint32 array the
 * kernels read): 1/8 + 3 (+1) bytes per position in all.  The ids of the array must then be below n_reads + n_refs, as
 * the file format demands (Tools.h:48); the environment variable LIME_PACK_DA=0 leaves da as it is.
 */
int  lime_set_alpha(lime_ctx *ctx, uint32_t alpha);
/* The host half on its own (no CUDA): out[i] = min(lcp[i], 255) with `threads` host threads (<= 0:
 * lime_host_threads()), and the same through the chunk loop lime_load runs (slots stage buffers of
 * `chunk` positions) over a simulated in-order copy queue that performs each copy copy_delay_us after
 * taking it.  Exposed for the CPU tests and for callers that narrow while they read a file. */
int  lime_host_threads(void);     /* LIME_HOST_THREADS, else min(cores in the affinity mask, 16) */
int  lime_host_narrow_lcp(const uint32_t *lcp, uint8_t *out, uint64_t n, int threads);
int  lime_host_narrow_lcp_pipelined(const uint32_t *lcp, uint8_t *out, uint64_t n, uint64_t chunk,
                                    int threads, int slots, int copy_delay_us);
/* The same two for the bit transfer of lime_set_alpha: out = ceil(n / 8) bytes, bit j of out[i] = lcp[8 i + j] < alpha,
 * bits beyond n set; chunk must be a multiple of 512. */
int  lime_host_pack_lcp(const uint32_t *lcp, uint8_t *out, uint64_t n, uint32_t alpha, int threads);
int  lime_host_pack_lcp_pipelined(const uint32_t *lcp, uint8_t *out, uint64_t n, uint64_t chunk, uint32_t alpha,
                                  int threads, int slots, int copy_delay_us);
/* da as 3 bytes per position (out = 3 n bytes, little endian), alone and through the same chunk loop together with
 * the lcp bits. */
int  lime_host_pack_da3(const uint32_t *da, uint8_t *out, uint64_t n);
int  lime_host_pack_lcp_da3_pipelined(const uint32_t *lcp, const uint32_t *da, uint8_t *out_bits, uint8_t *out_da3, uint64_t n,
                                      uint64_t chunk, uint32_t alpha, int threads, int slots, int copy_delay_us);

/* ---- step 1 : ClusterLCP.cpp:139-283 ------------------------------------------------------ */

/* Detect the alpha-clusters owned by this shard.  n_clusters / max_len are the shard's values of
 * nClusters / maxLen (reduction(+) / reduction(max), ClusterLCP.cpp:140); clusters are kept on the
 * device in ascending pStart order (the order of the reference run with threads = 1). */
int  lime_cluster_detect(lime_ctx *ctx, uint32_t alpha, uint64_t *n_clusters, uint64_t *max_len);
/* Copy the cluster records ({pStart,len}, global positions) to host memory, e.g. to write .clrs. */
int  lime_clusters_copy(lime_ctx *ctx, lime_cluster *out, uint64_t cap);
/* Replace the resident cluster list by caller-supplied records (step 2 started from a .clrs file,
 * ClusterBWT_DA.cpp:641-654); any order.  pStart is a global position. */
int  lime_clusters_set(lime_ctx *ctx, const lime_cluster *cl, uint64_t n_clusters, uint64_t max_len);

/* ---- steps 1 + 2a fused ------------------------------------------------------------------- */

/*
 * One pass over lcp / da / ebwt that detects the alpha-clusters of this shard (as
 * lime_cluster_detect, ClusterLCP.cpp:139-283) AND accumulates their similarity (as
 * lime_similarity, ClusterBWT_DA.cpp:256-358), without materialising the cluster list: what
 * `LiME_paired.sh` does with ClusterLCP followed by clusterAnalyze, minus the .clrs round trip.
 * keep_clusters != 0 also keeps the records on the device for lime_clusters_copy (.clrs output).
 * n_clusters / max_len are the values ClusterLCP writes to <stem>.out; max_len > sizeMaxBuf is
 * the error of ClusterBWT_DA.cpp:558-562 (LIME_ERR_MAXLEN).  May be called once per load.
 */
int  lime_cluster_similarity(lime_ctx *ctx, uint32_t alpha, int ebwt_mode, int keep_clusters,
                             uint64_t *n_clusters, uint64_t *max_len);

/*
 * One strand of LiME_paired.sh (:44-68: ClusterLCP, then ClusterBWT_DA) without a host synchronisation:
 * lime_cluster_similarity + lime_choose(slot, norm, beta, all reads) queued on the ctx stream -- fused kernel,
 * block path, one choose pass that also proves the absence of uint8 carries (byte sum) and leaves the matrix
 * zero for the next lime_load.  Errors of the strand (halo, maxLen > sizeMaxBuf) and its scalars surface at the
 * next lime_classify, which returns LIME_ERR_RETRY when the strand has to be redone with the synchronising
 * calls (record buffer sized from the previous strands too small; a cell above 255).  The first strand of a
 * context runs synchronously (it sizes the buffers).  lime_strand_status: nClusters, maxLen (the values of
 * <stem>.out), records incl. sentinel, updates -- valid once the slot has been read back by lime_classify.
 */
int  lime_strand_run(lime_ctx *ctx, int slot, uint32_t alpha, int ebwt_mode, uint32_t norm, float beta);
int  lime_strand_status(lime_ctx *ctx, int slot, uint64_t out[4]);

/* ---- step 2 : ClusterBWT_DA.cpp:256-358, 361-450 ------------------------------------------ */

/* Accumulate t(read, genome) of every resident cluster into the similarity matrix
 * (SimArray_, ClusterBWT_DA.cpp:606-611).  ebwt_mode 1 = EBWT=1 scoring (:107-190), 0 = EBWT=0
 * (:192-252).  Cells are exact modulo 256 (uint8 wrap, Tools.h:68).  May be called once per load. */
int  lime_similarity(lime_ctx *ctx, int ebwt_mode);

/* Bind caller-owned device memory as the accumulator: rows x pitch bytes, pitch a multiple of 16
 * and >= n_refs, rows >= n_reads; one uint8 cell per (read, genome).  Must be called before
 * lime_load; used by the multi-GPU host to reduce-scatter partial matrices (uint8 sum wraps like
 * the reference's cell).  Pass NULL to return to the internal allocation. */
int  lime_sim_bind(lime_ctx *ctx, void *dev_matrix, uint64_t rows, uint64_t pitch);
/* Row pitch in bytes the ctx uses for n_refs genomes. */
uint64_t lime_sim_pitch(uint32_t n_refs);
/* Dense copy of the matrix, [n_reads][n_refs] uint8 (the SMALL=1 dump, ClusterBWT_DA.cpp:672-681); in exchange
 * mode (lime_xchg_*) the selected slab, [reads_per_rank][n_refs]. */
int  lime_sim_copy(lime_ctx *ctx, uint8_t *out);

/*
 * clusterChoose for reads [read_begin, read_end): thresholds each row with beta and builds the
 * images of <fasta>.res.bin / .res.pos for that read range in device memory, stored in strand
 * slot `slot` (0..3 = mate1 F, mate1 RC, mate2 F, mate2 RC, the argv order of Classify).
 * Record 0 is the sentinel; pos values index into this range's own record list.
 * norm = readLen + 1 - alpha (ClusterBWT_DA.cpp:555).  n_records includes the sentinel.
 */
int  lime_choose(lime_ctx *ctx, int slot, uint32_t norm, float beta,
                 uint32_t read_begin, uint32_t read_end, uint64_t *n_records);
int  lime_choose_copy(lime_ctx *ctx, int slot, lime_pair_sim *bin, uint64_t *pos);
/* Load a strand slot from host images of .res.bin / .res.pos (step 3 started from files). */
int  lime_slot_set(lime_ctx *ctx, int slot, const lime_pair_sim *bin, uint64_t n_records,
                   const uint64_t *pos, uint32_t n_reads_in_slot);

/* ---- multi-GPU exchange of partial similarity ------------------------------------------------ */

/*
 * With the position range sharded over GPUs every GPU accumulates a PARTIAL similarity matrix over
 * all reads.  Reads are owned by contiguous ranges; before lime_choose the owner needs the sum.  This is the
 * exact route (a cell above 255 wraps as the reference's does) behind the faster lime_xchg_* calls below.
 * lime_sim_export writes the non-zero cells of rows [read_begin, read_end) as triples
 * {uint32 read, uint32 genome << 8 | count} (ascending read, then genome) to DEVICE memory
 * `dev_triples` (capacity `cap` triples, 8 bytes each) -- what is sent to the owner of that range;
 * *n_triples is the number found (LIME_ERR_NOMEM if it exceeds cap: nothing past cap is written).
 * lime_sim_import adds received triples (DEVICE memory) into this GPU's matrix, uint8 wrap as the
 * reference's cell (Tools.h:68); (read, genome) keys must be unique within one call.
 */
int  lime_sim_export(lime_ctx *ctx, uint32_t read_begin, uint32_t read_end, void *dev_triples,
                     uint64_t cap, uint64_t *n_triples);
int  lime_sim_import(lime_ctx *ctx, const void *dev_triples, uint64_t n_triples);

/* ---- multi-GPU: similarity rows on the read owner, updates routed as records ------------------ */

/*
 * The route of north_star ("partial (read, genome, count) triples are routed to their owning GPU by read id")
 * without a partial matrix over all reads: rank k owns reads [k * reads_per_rank, (k+1) * reads_per_rank) and
 * keeps only their rows (one slab of reads_per_rank x pitch uint8 per strand) plus an inbox with one region of
 * inbox_cap records per source rank and strand, which the other processes map (CUDA IPC).  The reference's
 * `SimArray_[read][genome] += t` (ClusterBWT_DA.cpp:183,247) becomes a record {read, genome << 8 | t}:
 *   lime_xchg_accumulate  lime_cluster_similarity with every update appended to a local list (capacity
 *                         flat_cap records; two lists, used alternately by the strands), then -- on a side
 *                         stream of the library -- sorted by owner and written into this rank's region of each
 *                         owner's inbox: bulk stores over NVLink / NVSwitch, nothing remote is read or updated
 *                         atomically.  Inside an owner the records are also split by row block of its slab (each
 *                         block small enough for L2).  Behind the records the rank stores its record counts and
 *                         a flag (the number of this accumulation) into a mailbox at the head of the owner's inbox;
 *   lime_xchg_apply       (side stream) waits on the device until every source's flag is up -- bounded, 20 s --
 *                         then adds the records of all regions of this rank's inbox to the slab, row block after
 *                         row block (the slab's lines stay in L2 while their records arrive);
 *   lime_xchg_choose      (side stream) clusterChoose (ClusterBWT_DA.cpp:361-450) of the owned reads into strand
 *                         slot `strand`.
 * None of the three synchronises with the host and there is no collective: records, counts and flags travel as
 * stores to peer memory, and the side-stream work of strand s runs beside lime_load + lime_xchg_accumulate of
 * strand s + 1 on the context's stream.  lime_classify joins the two streams, classifies the owned reads and reads
 * the strands' device state back: LIME_ERR_RETRY means the job must be redone through lime_sim_export /
 * lime_sim_import (a cell exceeded 255, where the reference wraps, or a list / inbox was too small); the ranks
 * agree on that by their own collective -- which every job needs anyway before the next one starts, because a job
 * rewrites the inboxes of the one before.  lime_strand_status gives nClusters / maxLen / records / updates of
 * this rank's shard.
 * world, reads_per_rank, n_refs, n_strands and inbox_cap of lime_xchg_alloc must be the same on every rank: a source
 * addresses its region inside the owner's inbox with them (flat_cap is local).  Every rank must make the same
 * sequence of lime_xchg_accumulate calls (the flag value is their count per strand).
 * Call order on every rank: lime_xchg_alloc -> all-gather the 64-byte handles -> lime_xchg_open (once) -> barrier;
 * per job and strand: lime_load, lime_xchg_accumulate, lime_xchg_apply, lime_xchg_choose; lime_classify.
 * lime_xchg_open(ctx, NULL) is a loopback for tests on one GPU: all owners' regions are local, rank k applies
 * region k.  lime_xchg_status synchronises and returns nClusters, maxLen, updates, list slots used, sum of t
 * applied, byte sum of the slab, records, error bits of one strand.
 */
#define LIME_IPC_HANDLE_BYTES 64
#define LIME_XCHG_MAX         16
int  lime_xchg_alloc(lime_ctx *ctx, int world, int rank, uint32_t reads_per_rank, uint32_t n_refs, int n_strands,
                     uint64_t flat_cap, uint64_t inbox_cap, uint8_t *ipc_handle);
int  lime_xchg_open(lime_ctx *ctx, const uint8_t *ipc_handles);
int  lime_xchg_close(lime_ctx *ctx);
int  lime_xchg_accumulate(lime_ctx *ctx, uint32_t alpha, int ebwt_mode, int strand);
int  lime_xchg_apply(lime_ctx *ctx, int strand);
int  lime_xchg_choose(lime_ctx *ctx, int strand, uint32_t norm, float beta, uint32_t n_own_reads);
int  lime_xchg_status(lime_ctx *ctx, int strand, uint64_t out[8]);

/* ---- clusterChoose without any allocation -------------------------------------------------------------- */

/*
 * No counterpart in the reference: added for callers that cannot call
 * cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  It is lime_choose
 * (clusterChoose, ClusterBWT_DA.cpp:361-450) of all rows of a caller-owned matrix, context-free, on the current
 * device and the given stream, and it calls no cudaMalloc / cudaFree or other allocating routine: the kernel's
 * scratch (tile counter, record count, one look-back word per 256 reads) is a static device array of
 * LIME_CHOOSE_STATIC_SCRATCH_BYTES that every GPU loading the library holds until exit.
 *   dev_sim        DEVICE, [n_reads][pitch] uint8, 16-byte aligned, pitch a multiple of 16 and >= n_refs; read only
 *   dev_bin        DEVICE, bin_cap records: the .res.bin image, record 0 = the sentinel (writes are guarded by bin_cap;
 *                  1 + n_reads + non-zero cells always suffices)
 *   dev_pos        DEVICE, n_reads uint64: the .res.pos image
 *   dev_n_records  DEVICE uint64 or NULL: the record count incl. the sentinel, written in stream order (no
 *                  synchronisation when n_records is NULL)
 *   n_records      HOST uint64 or NULL: the same, after a stream synchronisation; then LIME_ERR_NOMEM if it exceeds
 *                  bin_cap.  One of the two must be given.
 *   err            optional buffer for the message of a failure (err_len bytes)
 * n_reads above LIME_CHOOSE_STATIC_MAX_READS is rejected with LIME_ERR_ARG.  All calls share the scratch array, which
 * keeps the last call's contents: every call first clears the part it uses (in stream order), and calls that overlap
 * on one GPU (two streams, two threads) race -- serialise them.
 */
#define LIME_CHOOSE_STATIC_MAX_READS      16777216ull
#define LIME_CHOOSE_STATIC_SCRATCH_BYTES  (256ull + (LIME_CHOOSE_STATIC_MAX_READS / 256ull) * 8ull)
int  lime_choose_static(uint8_t *dev_sim, uint64_t pitch, uint32_t n_reads, uint32_t n_refs, uint32_t norm, float beta,
                        lime_pair_sim *dev_bin, uint64_t bin_cap, uint64_t *dev_pos, uint64_t *dev_n_records,
                        uint64_t *n_records, void *stream, char *err, uint64_t err_len);

/* ---- step 3 : Classify.cpp:474-702 -------------------------------------------------------- */

/*
 * Classify the reads of slots 0..n_files-1 (n_files = 2 or 4, Classify.cpp:334-338).
 *   v_corRef  [n_refs]     genome -> taxid at tax_rank (FixRank, Classify.cpp:33-89)
 *   m_corRef  [6][n_refs]  genome -> taxid per rank, or NULL      (HIGHER=1, :61-67)
 *   higher    0/1          the HIGHER build variant (:262-302, 672-677)
 * Outputs (host, one per read of the slot range; NULL: not copied): type 'C'/'U'/'A'/'H', taxid, sim
 * (typeOutput, Tools.h:100-104; Assign, Classify.cpp:157-162).
 */
int  lime_classify(lime_ctx *ctx, int n_files, uint32_t n_refs,
                   const uint32_t *v_corRef, const uint32_t *m_corRef, int tax_rank, int higher,
                   uint8_t *out_type, uint32_t *out_tax, float *out_sim);

/* The results of the last lime_classify again, into host (LIME_MEM_HOST) or device (LIME_MEM_DEVICE)
 * buffers of the caller; NULL pointers are skipped.  With lime_classify(..., NULL, NULL, NULL) this
 * keeps the assignments on the device, e.g. for a gather over NVLink in a multi-GPU job. */
int  lime_classify_result(lime_ctx *ctx, uint8_t *type, uint32_t *tax, float *sim, int mem);

/* ---- instrumentation ---------------------------------------------------------------------- */

/* Kernels launched by this ctx since creation (bench.py's gpu_launches). */
uint64_t lime_launch_count(const lime_ctx *ctx);
/* When enabled, every kernel launch is bracketed by CUDA events on the ctx stream. */
int  lime_prof_enable(lime_ctx *ctx, int on);
/* Synchronise and report: for kernel family k (0..lime_prof_kernels()-1) its name, launch count
 * and total device milliseconds since the last lime_prof_reset. */
int  lime_prof_kernels(void);
int  lime_prof_get(lime_ctx *ctx, int k, const char **name, uint64_t *launches, double *ms);
int  lime_prof_reset(lime_ctx *ctx);
/* Counters of the last lime_similarity call: [0] clusters taken by the warp path, [1] clusters
 * taken by the block path, [2] (read,genome) updates issued, [3] 1 if the packed uint8 pass
 * detected a carry and the exact 32-bit pass was run. */
int  lime_sim_stats(lime_ctx *ctx, uint64_t out[4]);

#ifdef __cplusplus
}
#endif
#endif /* LIME_CUDA_H_ */
