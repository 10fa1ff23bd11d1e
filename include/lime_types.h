/*
 * lime_types.h -- on-disk / in-memory type widths of the LiME classification path.
 *
 * These are the widths and record layouts the reference fixes in src/Tools.h and
 * that every file format of the path depends on:
 *
 *   ebwt symbol      uint8_t   (Tools.h:47  dataTypedimAlpha)
 *   lcp value        uint32_t  (Tools.h:48  dataTypelenSeq)
 *   document id      uint32_t  (Tools.h:49,55-59  dataTypeNSeq, dataTypeNumSeq==0)
 *   position         uint64_t  (Tools.h:50,61-65  dataTypeNChar, dataTypeNumChar==1)
 *   similarity cell  uint8_t   (Tools.h:51,67-77  dataTypeSim, wraps mod 256)
 *
 * Plain C, no CUDA / torch types.  Included by the C ABI (lime_cuda.h), the
 * CUDA translation units, the host executables and the oracle.
 */
#ifndef LIME_TYPES_H_
#define LIME_TYPES_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define LIME_ALF            16        /* symbol classes, Tools.h:39 (ALF)            */
#define LIME_SIZE_MAX_BUF   65536     /* max cluster length step 2 accepts, Tools.h:35 */
#define LIME_USIM_MAX       255       /* Tools.h:76 (USim_MAX for uchar cells)       */
#define LIME_RANKS          6         /* Classify.cpp:24 (RANK)                      */
#define LIME_MAXFILES       4         /* Classify.cpp:23 (MAXFILES)                  */
#define LIME_ERROR          0.02      /* Tools.h:37 (ERROR); cast to float at each use */

/* Tools.h:85-88 ElementCluster: one alpha-cluster = positions [pStart, pStart+len) */
typedef struct lime_cluster {
    uint64_t pStart;
    uint64_t len;
} lime_cluster;

/* Tools.h:95-98 pair_sim: one record of <fasta>.res.bin */
typedef struct lime_pair_sim {
    float    sim;
    uint32_t idRef;
} lime_pair_sim;

/* <stem>.out, 28 bytes, written field by field without padding (ClusterLCP.cpp:304-308) */
typedef struct lime_aux {
    uint32_t numReads;
    uint32_t numGenomes;
    uint32_t alpha;
    uint64_t maxLen;
    uint64_t nClusters;
} lime_aux;
#define LIME_AUX_BYTES 28

/* per-read result of step 3 (Tools.h:100-104 typeOutput) */
typedef struct lime_assign {
    uint32_t taxid;
    float    sim;
    uint8_t  type;   /* 'C','U','A','H' */
} lime_assign;

#ifdef __cplusplus
}
#endif
#endif /* LIME_TYPES_H_ */
