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

#ifdef __cplusplus
}
#endif
#endif
