/*
 * lime_oracle.c -- CPU restatement of LiME's classification hot path.
 *
 * TEST INFRASTRUCTURE ONLY.  This file is the checker the CUDA path is compared
 * against; nothing in lime_b200/ may call it.  Only tests/, __graft_entry__.smoke()
 * and bench.py's cpu_baseline leg load the shared library built from it.
 *
 * It restates, in plain sequential C, the algorithm of the reference executables
 *   ClusterLCP     (/root/reference/src/ClusterLCP.cpp)
 *   ClusterBWT_DA  (/root/reference/src/ClusterBWT_DA.cpp, EBWT=1 and EBWT=0, BIN=1)
 *   Classify       (/root/reference/src/Classify.cpp, HIGHER=0 and HIGHER=1, BIN=1)
 * Each function cites the reference lines it follows.
 *
 * Parity pin: the reference ships no golden vectors or tests, so the
 * oracle is pinned against the reference itself: oracle/Makefile compiles the untouched
 * reference sources into oracle/_ref/ and tests/test_oracle_vs_reference.py diffs
 * the files both produce (.clrs, .out, .res.bin, .res.pos, CSV) on the corpus of
 * tests/corpus.py; tests/golden/ holds reference outputs generated that way (with
 * the generating script) so the pin also holds where /root/reference is absent.
 */
#include <stdint.h>
#include <stdlib.h>
#include <stdio.h>
#include <string.h>
#include "../include/lime_types.h"

/* ------------------------------------------------------------------------- */
/* Step 1: alpha-cluster detection                                            */
/* ------------------------------------------------------------------------- */

/*
 * Single logical pass equal to ClusterLCP.cpp:186-272 for any thread count (the
 * threads only re-partition the same scan, :150-202,244-264).
 *
 *  - the scan skips the leading run of lcp >= alpha (:196-202); positions before
 *    the first lcp < alpha are never clustered;
 *  - lcp[i] >= alpha opens (pStart = i-1, :16-24) or extends a cluster and
 *    records whether da[i-1] / da[i] is a read (< n_reads) or a genome (:20-30);
 *  - lcp[i] < alpha closes it; it is emitted only if it saw a read and a
 *    genome (:220-221), len = i - pStart (:36);
 *  - an open cluster at end of input is closed by the last thread (:244-245).
 * Returns the number of clusters written to out (ascending pStart); *max_len is
 * the longest emitted cluster (:37-38).  out may be NULL to count only.
 */
uint64_t oracle_cluster_detect(const uint32_t *lcp, const uint32_t *da, uint64_t n,
                               uint32_t n_reads, uint32_t alpha,
                               lime_cluster *out, uint64_t out_cap, uint64_t *max_len)
{
    uint64_t count = 0, longest = 0;
    uint64_t p = 0;
    while (p < n && lcp[p] >= alpha) p++;
    int open = 0, has_r = 0, has_g = 0;
    uint64_t start = 0;
    for (uint64_t i = p + 1; i < n; i++) {
        if (lcp[i] >= alpha) {
            if (!open) {
                open = 1; start = i - 1;
                if (da[i - 1] < n_reads) has_r = 1; else has_g = 1;
            }
            if (!(has_r && has_g)) {
                if (da[i] < n_reads) has_r = 1; else has_g = 1;
            }
        } else {
            if (open && has_r && has_g) {
                uint64_t len = i - start;
                if (len > longest) longest = len;
                if (out && count < out_cap) { out[count].pStart = start; out[count].len = len; }
                count++;
            }
            open = 0; has_r = 0; has_g = 0;
        }
    }
    if (open && has_r && has_g) {
        uint64_t len = n - start;
        if (len > longest) longest = len;
        if (out && count < out_cap) { out[count].pStart = start; out[count].len = len; }
        count++;
    }
    if (max_len) *max_len = longest;
    return count;
}

/* ------------------------------------------------------------------------- */
/* Step 2: similarity accumulation                                            */
/* ------------------------------------------------------------------------- */

/* ClusterBWT_DA.cpp:455-470 -- symbol classes; every byte not listed maps to
 * class 0 because unordered_map::operator[] default-inserts 0 (:96,123). */
static uint8_t class_of[256];
/* ClusterBWT_DA.cpp:472-487 -- class a (4..14) is compatible with base i (0..3) */
static const uint8_t compat[15][4] = {
    {1,0,0,0}, {0,1,0,0}, {0,0,1,0}, {0,0,0,1},
    {1,0,1,0}, {0,1,0,1}, {0,1,1,0}, {1,0,0,1}, {0,0,1,1}, {1,1,0,0},
    {0,1,1,1}, {1,0,1,1}, {1,1,0,1}, {1,1,1,0}, {1,1,1,1}
};
static int tables_ready = 0;
static void init_tables(void)
{
    if (tables_ready) return;
    memset(class_of, 0, sizeof class_of);
    const char *sym = "ACGTRYSWKMBDHVN";
    for (int c = 0; c < 15; c++) class_of[(uint8_t)sym[c]] = (uint8_t)c;
    class_of[0] = 15;
    tables_ready = 1;
}

void oracle_class_table(uint8_t *dst256) { init_tables(); memcpy(dst256, class_of, 256); }

/*
 * Score of one (read, genome) pair inside one cluster, EBWT=1
 * (ClusterBWT_DA.cpp:131-177).  All quantities are uint8 and wrap.  R = read's
 * per-class counts (wrapping, :123), G = genome's per-class counts (saturated
 * at 255, :96-97).  Equal classes are matched first (:133-144); leftovers are
 * cross-matched between bases i=0..3 and IUPAC classes a=4..14 (:146-177).  In
 * the "genome IUPAC vs read base" branch the reference zeroes one operand
 * before subtracting it from the other (:153-154,158-159), so neither leftover
 * is reduced -- reproduced here by simply not decrementing.
 */
uint8_t oracle_score_ebwt(const uint8_t R[16], const uint8_t G[16])
{
    uint8_t t = 0, rr[16], gg[16];
    for (int c = 0; c < LIME_ALF; c++) {
        if (R[c] > G[c]) { t = (uint8_t)(t + G[c]); rr[c] = (uint8_t)(R[c] - G[c]); gg[c] = 0; }
        else             { t = (uint8_t)(t + R[c]); gg[c] = (uint8_t)(G[c] - R[c]); rr[c] = 0; }
    }
    for (int i = 0; i < 4; i++) {
        for (int a = 4; a < LIME_ALF - 1; a++) {
            if (!compat[a][i]) continue;
            if (gg[a] > 0) {
                if (gg[a] > rr[i]) { t = (uint8_t)(t + rr[i]); rr[i] = 0; }
                else               { t = (uint8_t)(t + gg[a]); gg[a] = 0; }
            }
            if (rr[a] > 0) {
                if (rr[a] > gg[i]) { t = (uint8_t)(t + gg[i]); rr[a] = (uint8_t)(rr[a] - gg[i]); gg[i] = 0; }
                else               { t = (uint8_t)(t + rr[a]); gg[i] = (uint8_t)(gg[i] - rr[a]); rr[a] = 0; }
            }
        }
    }
    return t;
}

static int cmp_u64(const void *a, const void *b)
{
    uint64_t x = *(const uint64_t *)a, y = *(const uint64_t *)b;
    return (x > y) - (x < y);
}

/*
 * clusterAnalyze (ClusterBWT_DA.cpp:256-358) over a cluster list.  For each
 * cluster the (da, class) pairs are sorted by da then class (:316-322), split
 * at the first genome (:329,336), genomes are summarised first
 * (Update_ref_symb :81-105 / :213-228), then every read is scored against
 * every genome present (:107-190 / :232-250) and the uint8 cell is bumped
 * (wrapping, :183,247).  ebwt == NULL selects the EBWT=0 variant in which only
 * occurrence counts are compared (t = min(#read mod 256, min(#genome,255))).
 * sim is a dense row-major [n_reads][n_refs] uint8 matrix the caller zeroed.
 * Returns 0, or -1 if a cluster is longer than sizeMaxBuf (:558-562).
 */
int oracle_similarity(const lime_cluster *cl, uint64_t n_cl,
                      const uint32_t *da, const uint8_t *ebwt,
                      uint32_t n_reads, uint32_t n_refs, uint8_t *sim)
{
    init_tables();
    uint64_t *keys = (uint64_t *)malloc(sizeof(uint64_t) * LIME_SIZE_MAX_BUF);
    uint32_t *gid = (uint32_t *)malloc(sizeof(uint32_t) * LIME_SIZE_MAX_BUF);
    uint8_t (*gcnt)[16] = (uint8_t (*)[16])malloc(16 * (size_t)LIME_SIZE_MAX_BUF);
    if (!keys || !gid || !gcnt) { free(keys); free(gid); free(gcnt); return -2; }
    int rc = 0;
    for (uint64_t c = 0; c < n_cl; c++) {
        uint64_t s = cl[c].pStart, len = cl[c].len;
        if (len > LIME_SIZE_MAX_BUF) { rc = -1; break; }
        for (uint64_t e = 0; e < len; e++) {
            uint8_t k = ebwt ? class_of[ebwt[s + e]] : 0;
            keys[e] = ((uint64_t)da[s + e] << 8) | k;
        }
        qsort(keys, (size_t)len, sizeof(uint64_t), cmp_u64);
        uint64_t split = 0;
        while (split < len && (uint32_t)(keys[split] >> 8) < n_reads) split++;
        /* genomes present in the cluster with their saturated class counts */
        uint64_t ng = 0;
        for (uint64_t e = split; e < len;) {
            uint32_t id = (uint32_t)(keys[e] >> 8);
            memset(gcnt[ng], 0, 16);
            while (e < len && (uint32_t)(keys[e] >> 8) == id) {
                uint8_t k = (uint8_t)(keys[e] & 0xff);
                if (gcnt[ng][k] < LIME_USIM_MAX) gcnt[ng][k]++;
                e++;
            }
            gid[ng++] = id - n_reads;
        }
        /* each read against each genome */
        for (uint64_t e = 0; e < split;) {
            uint32_t id = (uint32_t)(keys[e] >> 8);
            uint8_t R[16];
            memset(R, 0, 16);
            while (e < split && (uint32_t)(keys[e] >> 8) == id) {
                R[keys[e] & 0xff]++;                     /* uint8, wraps */
                e++;
            }
            for (uint64_t j = 0; j < ng; j++) {
                uint8_t t;
                if (ebwt) t = oracle_score_ebwt(R, gcnt[j]);
                else      t = R[0] > gcnt[j][0] ? gcnt[j][0] : R[0];
                if (t > 0) {
                    uint8_t *cell = &sim[(size_t)id * n_refs + gid[j]];
                    *cell = (uint8_t)(*cell + t);
                }
            }
        }
    }
    free(keys); free(gid); free(gcnt);
    return rc;
}

/*
 * clusterChoose, BIN=1 (ClusterBWT_DA.cpp:361-450).  Record 0 is the {0.0f,0}
 * sentinel (:376-382).  For each read: maxSim over the row (:390-402),
 * normSim = (float)maxSim / norm (:404, norm is a uint converted to float);
 * if normSim > beta the header {normSim, count} and count x {(float)sim/norm,
 * genome} (ascending genome, non-zero cells only) are appended and pos[read] =
 * index of the header, else pos[read] = 0 (:405-436).
 * Returns the number of records (including the sentinel); bin may be NULL to
 * size the output first.
 */
uint64_t oracle_choose(const uint8_t *sim, uint32_t n_reads, uint32_t n_refs,
                       uint32_t norm, float beta, lime_pair_sim *bin, uint64_t *pos)
{
    uint64_t total = 0;
    if (bin) { bin[0].sim = 0.0f; bin[0].idRef = 0; }
    total = 1;
    for (uint32_t r = 0; r < n_reads; r++) {
        const uint8_t *row = sim + (size_t)r * n_refs;
        uint8_t max_sim = 0;
        uint32_t nnz = 0;
        for (uint32_t j = 0; j < n_refs; j++) {
            if (row[j] > max_sim) max_sim = row[j];
            if (row[j] > 0) nnz++;
        }
        float norm_sim = (float)max_sim / (float)norm;
        if (norm_sim > beta) {
            if (pos) pos[r] = total;
            if (bin) { bin[total].sim = norm_sim; bin[total].idRef = nnz; }
            total++;
            for (uint32_t j = 0; j < n_refs; j++) {
                if (row[j] == 0) continue;
                if (bin) { bin[total].sim = (float)row[j] / (float)norm; bin[total].idRef = j; }
                total++;
            }
        } else if (pos) {
            pos[r] = 0;
        }
    }
    return total;
}

/* ------------------------------------------------------------------------- */
/* Step 3: classification                                                     */
/* ------------------------------------------------------------------------- */

/*
 * FixRank (Classify.cpp:33-89).  The reference tokenises the whole stream with
 * getline(';') x6 + getline('\n') and pushes a row only while the stream is
 * still good() after reading it, so a last row without trailing newline is
 * dropped; CRLF and a trailing blank line are harmless (atoi stops at '\r').
 * A row whose tax_rank column is empty is skipped for v (:59-60) but still
 * advances the HIGHER matrix row (:61-66).  tax_rank == 0 maps genome -> its
 * own index (:73-88).  m (may be NULL) is [6][n_refs], zero-initialised here
 * (:384-389); m[i][g] = column i+1 for i >= tax_rank-1.
 * Returns the number of v entries produced (caller compares with n_refs,
 * :395-398), or -1 if the file cannot be opened.
 */
typedef struct { const char *p, *end; int eof, fail; } tok_t;

static void tok_field(tok_t *t, char delim, char *dst, size_t cap)
{
    if (t->eof || t->fail) { t->fail = 1; return; }
    size_t n = 0; int extracted = 0;
    dst[0] = 0;
    for (;;) {
        if (t->p >= t->end) { t->eof = 1; break; }
        char ch = *t->p++;
        extracted = 1;
        if (ch == delim) break;
        if (n + 1 < cap) dst[n++] = ch;
    }
    dst[n] = 0;
    if (!extracted) t->fail = 1;
}

static void tok_row(tok_t *t, char lin[7][256])
{
    for (int s = 0; s < LIME_RANKS; s++) tok_field(t, ';', lin[s], 256);
    tok_field(t, '\n', lin[6], 256);
}

int64_t oracle_fix_rank(const char *path, int tax_rank, uint32_t n_refs,
                        uint32_t *v, uint64_t v_cap, uint32_t *m)
{
    FILE *f = fopen(path, "rb");
    if (!f) return -1;
    fseek(f, 0, SEEK_END);
    long sz = ftell(f);
    fseek(f, 0, SEEK_SET);
    char *buf = (char *)malloc((size_t)sz + 1);
    if (fread(buf, 1, (size_t)sz, f) != (size_t)sz) { fclose(f); free(buf); return -1; }
    fclose(f);
    tok_t t = { buf, buf + sz, 0, 0 };
    char lin[7][256];
    if (m) memset(m, 0, sizeof(uint32_t) * LIME_RANKS * (size_t)n_refs);
    tok_row(&t, lin);                 /* header, :48-50 */
    tok_row(&t, lin);                 /* first data row, :53-55 */
    uint64_t n = 0, nv = 0;
    while (!t.eof && !t.fail) {
        if (tax_rank > 0) {
            if (lin[tax_rank][0] != 0) { if (nv < v_cap) v[nv] = (uint32_t)atoi(lin[tax_rank]); nv++; }
        } else {
            if (nv < v_cap) v[nv] = (uint32_t)n;
            nv++;
        }
        if (m && n < n_refs) {
            for (int i = tax_rank - 1; i < LIME_RANKS; i++) {
                if (i < 0) continue;  /* tax_rank 0 with HIGHER indexes m[-1] in the reference: unsupported */
                if (lin[i + 1][0] != 0) m[(size_t)i * n_refs + n] = (uint32_t)atoi(lin[i + 1]);
            }
        }
        n++;
        tok_row(&t, lin);
    }
    free(buf);
    return (int64_t)nv;
}

typedef struct { float max; uint32_t size; const lime_pair_sim *e; } row_t;

static int taxa_single(const uint32_t *ids, uint32_t n, const uint32_t *map, uint32_t *tax)
{
    for (uint32_t i = 1; i < n; i++) if (map[ids[i]] != map[ids[0]]) return 0;
    *tax = map[ids[0]];
    return n > 0;
}

/*
 * Per-read decision of Classify.cpp:483-687 (BIN=1).  rows[i] is strand file i
 * (0 = mate1 F, 1 = mate1 RC, 2 = mate2 F, 3 = mate2 RC; pairs 0+3 and 1+2,
 * :221-224,600-603).  scratch = n_files*n_refs floats, ids = n_refs uint32.
 */
static void classify_read(int nf, const row_t *rows, uint32_t n_refs,
                          const uint32_t *v, const uint32_t *m, int tax_rank, int higher,
                          float *scratch, uint32_t *ids, float *cg,
                          uint8_t *o_type, uint32_t *o_tax, float *o_sim)
{
    const float ERR = (float)LIME_ERROR;
    *o_type = 'U'; *o_tax = 0; *o_sim = 0.0f;
    /* :497-518 -- unclassified if every strand's header max is 0 */
    int any = 0; float highest = 0.0f;
    for (int i = 0; i < nf; i++)
        if (rows[i].max != 0.0f) { if (!any || rows[i].max > highest) highest = rows[i].max; any = 1; }
    if (!any) return;
    /* stage A, :520-556 -- strands within ERR of the best max; their near-max genomes */
    uint32_t n_i = 0;
    for (int i = 0; i < nf; i++) {
        if (rows[i].max == 0.0f) continue;
        float d = highest - rows[i].max;
        if (!(d < ERR)) continue;
        for (uint32_t j = 0; j < rows[i].size; j++) {
            float dd = rows[i].max - rows[i].e[j].sim;      /* :113,120 */
            if (dd < ERR) {
                uint32_t g = rows[i].e[j].idRef, k;
                for (k = 0; k < n_i; k++) if (ids[k] == g) break;
                if (k == n_i) ids[n_i++] = g;
            }
        }
    }
    uint32_t tax;
    if (taxa_single(ids, n_i, v, &tax)) { *o_type = 'C'; *o_tax = tax; *o_sim = highest; return; }
    /* stage B, :557-640 -- per-strand sims of the candidate genomes, mates summed */
    for (uint32_t e = 0; e < n_i; e++)
        for (int i = 0; i < nf; i++) {
            float s = 0.0f;
            for (uint32_t j = 0; j < rows[i].size; j++)
                if (rows[i].e[j].idRef == ids[e]) { s = rows[i].e[j].sim; break; }
            cg[(size_t)i * n_refs + e] = s;
        }
    float maxsum[2] = { 0.0f, 0.0f };
    for (int k = 0; k < 2; k++)
        for (uint32_t e = 0; e < n_i; e++) {
            if (k == 0 && nf == 4) {
                cg[0 * (size_t)n_refs + e] += cg[3 * (size_t)n_refs + e];
                cg[1 * (size_t)n_refs + e] += cg[2 * (size_t)n_refs + e];
            }
            if (maxsum[k] < cg[(size_t)k * n_refs + e]) maxsum[k] = cg[(size_t)k * n_refs + e];
        }
    int win = -1;
    { float a = maxsum[1] + ERR, b = maxsum[0] + ERR;
      if (maxsum[0] > a) win = 0; else if (maxsum[1] > b) win = 1; }
    if (win >= 0) {
        int first = 1, single = 1; uint32_t tx = 0;
        for (uint32_t e = 0; e < n_i; e++)
            if (cg[(size_t)win * n_refs + e] == maxsum[win]) {
                if (first) { tx = v[ids[e]]; first = 0; }
                else if (v[ids[e]] != tx) single = 0;
            }
        if (!first && single) { *o_type = 'C'; *o_tax = tx; *o_sim = maxsum[win]; return; }
    }
    /* stage C, Exam_2 :171-260 -- dense over all genomes */
    for (size_t x = 0; x < (size_t)nf * n_refs; x++) scratch[x] = 0.0f;
    for (int i = 0; i < nf; i++)
        for (uint32_t j = 0; j < rows[i].size; j++)
            scratch[(size_t)i * n_refs + rows[i].e[j].idRef] += rows[i].e[j].sim;   /* :197 */
    float hi[2] = { 0.0f, 0.0f };
    for (int k = 0; k < 2; k++)
        for (uint32_t g = 0; g < n_refs; g++) {
            if (k == 0 && nf == 4) {
                scratch[0 * (size_t)n_refs + g] += scratch[3 * (size_t)n_refs + g];
                scratch[1 * (size_t)n_refs + g] += scratch[2 * (size_t)n_refs + g];
            }
            if (hi[k] < scratch[(size_t)k * n_refs + g]) hi[k] = scratch[(size_t)k * n_refs + g];
        }
    float h; uint32_t n_g = 0;
    const float *S0 = scratch, *S1 = scratch + n_refs;
    if (hi[0] > hi[1])      { h = hi[0]; for (uint32_t g = 0; g < n_refs; g++) { float d = h - S0[g]; if (d < ERR) ids[n_g++] = g; } }
    else if (hi[0] < hi[1]) { h = hi[1]; for (uint32_t g = 0; g < n_refs; g++) { float d = h - S1[g]; if (d < ERR) ids[n_g++] = g; } }
    else { h = hi[0]; for (uint32_t g = 0; g < n_refs; g++) { float d0 = h - S0[g], d1 = h - S1[g]; if (d0 < ERR || d1 < ERR) ids[n_g++] = g; } }
    if (taxa_single(ids, n_g, v, &tax)) { *o_type = 'C'; *o_tax = tax; *o_sim = h; return; }   /* :652-670 */
    /* :671-686 -- ambiguous, or (HIGHER) first rank at which the set agrees on a non-zero taxon, :262-302 */
    *o_type = 'A';
    if (higher && m) {
        for (int idx = tax_rank - 1; idx < LIME_RANKS; idx++) {
            if (idx < 0) continue;
            uint32_t tx;
            if (taxa_single(ids, n_g, m + (size_t)idx * n_refs, &tx) && tx != 0) {
                *o_type = 'H'; *o_tax = tx; *o_sim = h; return;
            }
        }
    }
}

/*
 * Classify main loop (Classify.cpp:474-702).  bin[i]/pos[i] are the images of
 * <strand i>.res.bin / .res.pos.  Outputs one (type, taxid, sim) per read as in
 * vOutput (:457-463, Assign :157-162).
 */
int oracle_classify(int n_files, const lime_pair_sim *const *bin, const uint64_t *const *pos,
                    uint32_t n_reads, uint32_t n_refs,
                    const uint32_t *v, const uint32_t *m, int tax_rank, int higher,
                    uint8_t *o_type, uint32_t *o_tax, float *o_sim)
{
    if (n_files != 2 && n_files != 4) return -1;                 /* :334-338 */
    float *scratch = (float *)malloc(sizeof(float) * 4 * (size_t)(n_refs ? n_refs : 1));
    float *cg = (float *)malloc(sizeof(float) * 4 * (size_t)(n_refs ? n_refs : 1));
    uint32_t *ids = (uint32_t *)malloc(sizeof(uint32_t) * (size_t)(n_refs ? n_refs : 1));
    for (uint32_t r = 0; r < n_reads; r++) {
        row_t rows[LIME_MAXFILES];
        for (int i = 0; i < n_files; i++) {
            uint64_t p = pos[i][r];
            if (p) { rows[i].max = bin[i][p].sim; rows[i].size = bin[i][p].idRef; rows[i].e = bin[i] + p + 1; }
            else   { rows[i].max = 0.0f; rows[i].size = 0; rows[i].e = NULL; }
        }
        classify_read(n_files, rows, n_refs, v, m, tax_rank, higher, scratch, ids, cg,
                      &o_type[r], &o_tax[r], &o_sim[r]);
    }
    free(scratch); free(cg); free(ids);
    return 0;
}

/*
 * Output loop (Classify.cpp:455,722-734): default ostream<<float is %g with 6
 * significant digits.
 */
int oracle_write_csv(const char *path, uint32_t n_reads,
                     const uint8_t *type, const uint32_t *tax, const float *sim)
{
    FILE *f = fopen(path, "w");
    if (!f) return -1;
    fputs("C/U/A/H,IdSeqRead,TaxID,maxSim\n", f);
    for (uint32_t r = 0; r < n_reads; r++) {
        if (type[r] == 'U')      fprintf(f, "U,%u,NA,0\n", r);
        else if (type[r] == 'A') fprintf(f, "A,%u,NA,0\n", r);
        else                     fprintf(f, "%c,%u,%u,%g\n", type[r] == 'C' ? 'C' : 'H', r, tax[r], (double)sim[r]);
    }
    fclose(f);
    return 0;
}
