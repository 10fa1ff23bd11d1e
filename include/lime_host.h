/*
 * lime_host.h -- host-side (CPU, non-hot-path) helpers of the LiME path that stay on the host:
 * lineage parsing and CSV output of step 3.  Exported from liblime_b200.so with C linkage so
 * the executables and language bindings share one implementation.
 */
#ifndef LIME_HOST_H_
#define LIME_HOST_H_
#include <stdint.h>
#include "lime_types.h"
#ifdef __cplusplus
extern "C" {
#endif

/*
 * FixRank (Classify.cpp:33-89).  Parses the ';'-separated lineage file (header + one row per
 * genome, 7 columns).  v[g] = taxid of genome g at column tax_rank (row skipped when the
 * field is empty, :59-60), or g itself when tax_rank == 0 (:73-76).  m (optional, [6][n_refs],
 * zero-filled here as in :384-389) gets m[i][g] = column i+1 for i >= tax_rank-1 (:61-66).
 * A last row that is not newline-terminated is dropped, exactly like the reference's
 * `while (f.good())` loop.  Returns the number of v entries found (may exceed v_cap; the caller
 * must compare with n_refs, :395-398) or -1 if the file cannot be opened.
 */
int64_t lime_host_fix_rank(const char *path, int tax_rank, uint32_t n_refs,
                           uint32_t *v, uint64_t v_cap, uint32_t *m);

/* Output loop of Classify.cpp:455,722-734: header line, then one line per read; sim printed like
 * `ostream << float` (6 significant digits). Returns 0, or -1 if the file cannot be created. */
int lime_host_write_csv(const char *path, uint32_t n_reads,
                        const uint8_t *type, const uint32_t *tax, const float *sim);

/*
 * The formats either side of the path.
 *
 * lime_host_read_gesa: de-interleaves <fasta>.<numSeq>.gesa, the output of egsa -- 13-byte records
 * { u32 text = document id, u32 suff, u32 lcp, u8 bwt } -- into the three arrays of the path, what
 * EGSAtoBCR.cpp:72-91 does through files.  Returns the number of complete records (a truncated last
 * record is dropped like in the reference loop); only the first `cap` are stored, so a call with
 * cap = 0 sizes the arrays.  -1 if the file cannot be opened.
 */
int64_t lime_host_read_gesa(const char *path, uint8_t *ebwt, uint32_t *lcp, uint32_t *da, uint64_t cap);

/*
 * BIN=0 build of the reference: ClusterBWT_DA writes <fasta>.res.txt instead of .res.bin / .res.pos --
 * one line per read, empty if the read does not pass beta, else "%.5f" of the normalised maximum
 * followed by "\t%u\t%.5f" per (genome, similarity) (ClusterBWT_DA.cpp:414,424,440) -- and Classify
 * parses those tokens (Classify.cpp:128-153).  The writer takes the images lime_choose_copy returns;
 * the reader rebuilds images (with the 5-decimal values of the text) for lime_slot_set.
 * write: 0, -1 (cannot create), -2 (inconsistent images).  read: records incl. the sentinel (may exceed
 * bin_cap: call again with a larger buffer), -1 (cannot open), -2 (fewer than n_reads lines).
 */
int lime_host_write_res_txt(const char *path, const lime_pair_sim *bin, uint64_t n_records,
                            const uint64_t *pos, uint32_t n_reads);
int64_t lime_host_read_res_txt(const char *path, uint32_t n_reads, lime_pair_sim *bin, uint64_t bin_cap,
                               uint64_t *pos);

#ifdef __cplusplus
}
#endif
#endif
