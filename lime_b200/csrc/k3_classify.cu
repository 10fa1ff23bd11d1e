// k3_classify.cu -- step 3, per-read taxon assignment (replaces Classify.cpp:474-702, BIN=1,
// HIGHER=0/1).
//
// The reference loads each strand's sparse row into vectors, builds candidate sets with
// std::find, and in Exam_2 scatters the rows into dense float[4][numTarg] arrays (:171-260).
// Rows written by step 2 are strictly ascending in genome id, so all of that is a k-way merge:
// one thread per read walks its 2 or 4 rows in lock step and sees, per genome, the strand sims
// s[i] (0 if absent).  Three merge passes:
//   pass 1  stage A (:520-556): strands within ERROR of the best max, their near-max genomes
//           (ReadSimilarity :110-126) -> set I, single taxon?  Also maxsum[0/1] of stage B over
//           I (:597-607) and highest[0/1] of Exam_2 over all genomes (:219-228).
//   pass 2  stage B (:608-640): taxa of I's genomes whose pair sum equals the winning maxsum.
//   pass 3  stage C (:230-253, :652-686): genomes within ERROR of h; single taxon -> 'C', else
//           'A' or (HIGHER) the first rank on which they share a non-zero taxon -> 'H' (:262-302).
// Float semantics are the reference's: float32 adds / subtracts / compares, ERROR cast to float
// at each use.  A genome absent from every row has sum 0, so when h - 0 < ERROR every genome of
// the database is in the stage-C set; the host precomputes that case (all_* fields).
// Precondition (checked, err bit0): ids strictly ascending and < n_refs within a row.
#include "common.cuh"

namespace lime {
namespace cls {

constexpr uint32_t NO_G = 0xffffffffu;

template <int NF>
struct Rows {
    const lime_pair_sim *e[NF];
    uint32_t size[NF];
    float mx[NF];
};

// k-way merge cursor
template <int NF>
struct Merge {
    uint32_t idx[NF];
    __device__ __forceinline__ void reset()
    {
#pragma unroll
        for (int i = 0; i < NF; i++) idx[i] = 0;
    }
    // next distinct genome id; fills s[] (0 if absent) and pres bitmask; returns false at the end
    __device__ __forceinline__ bool next(const Rows<NF> &R, uint32_t &g, float s[NF], unsigned &pres,
                                         unsigned &bad)
    {
        uint32_t head[NF];
        float sim[NF];
        g = NO_G;
#pragma unroll
        for (int i = 0; i < NF; i++) {
            head[i] = NO_G; sim[i] = 0.0f;
            if (idx[i] < R.size[i]) {
                const uint2 rec = __ldg(reinterpret_cast<const uint2 *>(R.e[i] + idx[i]));
                head[i] = rec.y; sim[i] = __uint_as_float(rec.x);
            }
            g = min(g, head[i]);
        }
        if (g == NO_G) return false;
        pres = 0;
#pragma unroll
        for (int i = 0; i < NF; i++) {
            if (head[i] == g) {
                s[i] = sim[i]; pres |= 1u << i; idx[i]++;
                if (idx[i] < R.size[i]) {
                    const uint32_t nx = __ldg(&R.e[i][idx[i]].idRef);
                    if (nx <= g) bad = 1;
                }
            } else s[i] = 0.0f;
        }
        return true;
    }
};

// "all the same value" tracker
struct Uniq {
    uint32_t val; bool any, same;
    __device__ __forceinline__ void init() { val = 0; any = false; same = true; }
    __device__ __forceinline__ void add(uint32_t v)
    {
        if (!any) { val = v; any = true; } else if (v != val) same = false;
    }
    __device__ __forceinline__ bool single() const { return any && same; }
};

template <int NF>
__device__ __forceinline__ void pair_sums(const float s[NF], float &p0, float &p1)
{
    if (NF == 4) { p0 = __fadd_rn(s[0], s[3]); p1 = __fadd_rn(s[1], s[2]); }   // :221-224, :600-603
    else         { p0 = s[0]; p1 = s[1]; }
}

template <int NF>
__global__ void __launch_bounds__(128) classify_kernel(ClsArgs P)
{
    const uint32_t r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= P.n_reads) return;
    const float ERR = (float)LIME_ERROR;
    unsigned bad = 0;

    Rows<NF> R;
    bool any = false;
    float highest = 0.0f;
#pragma unroll
    for (int i = 0; i < NF; i++) {
        const uint64_t p = P.pos[i][r];
        const uint64_t n_rec = P.n_records_dev[i] ? min((uint64_t)*P.n_records_dev[i], P.bin_cap[i]) : P.n_records[i];
        if (p) {
            if (p >= n_rec) { bad = 1; R.mx[i] = 0.0f; R.size[i] = 0; R.e[i] = nullptr; continue; }
            const lime_pair_sim h = P.bin[i][p];
            R.mx[i] = h.sim; R.size[i] = h.idRef; R.e[i] = P.bin[i] + p + 1;
            if (p + 1 + h.idRef > n_rec) { bad = 1; R.size[i] = 0; }
        } else { R.mx[i] = 0.0f; R.size[i] = 0; R.e[i] = nullptr; }
        if (R.mx[i] != 0.0f) {                                   // :506-509
            if (!any || R.mx[i] > highest) highest = R.mx[i];
            any = true;
        }
    }
    uint8_t type = 'U'; uint32_t tax = 0; float sim = 0.0f;      // :457-463
    if (any) {
        unsigned top = 0;                                        // :529-533
#pragma unroll
        for (int i = 0; i < NF; i++)
            if (R.mx[i] != 0.0f && __fsub_rn(highest, R.mx[i]) < ERR) top |= 1u << i;

        Merge<NF> M;
        uint32_t g; float s[NF]; unsigned pres;
        // ---- pass 1 ----
        Uniq uA; uA.init();
        float maxsum0 = 0.0f, maxsum1 = 0.0f, hi0 = 0.0f, hi1 = 0.0f;
        M.reset();
        while (M.next(R, g, s, pres, bad)) {
            if (g >= P.n_refs) { bad = 1; continue; }
            float p0, p1;
            pair_sums<NF>(s, p0, p1);
            if (hi0 < p0) hi0 = p0;                              // :225-226
            if (hi1 < p1) hi1 = p1;
            bool inI = false;
#pragma unroll
            for (int i = 0; i < NF; i++)
                if (((top & pres) >> i) & 1u)
                    if (__fsub_rn(R.mx[i], s[i]) < ERR) inI = true;   // :120
            if (inI) {
                uA.add(P.v[g]);                                  // :546-547
                if (maxsum0 < p0) maxsum0 = p0;                  // :604-605
                if (maxsum1 < p1) maxsum1 = p1;
            }
        }
        bool done = false;
        if (uA.single()) { type = 'C'; tax = uA.val; sim = highest; done = true; }   // :548-556
        // ---- pass 2 : stage B ----
        if (!done) {
            int win = -1;
            if (maxsum0 > __fadd_rn(maxsum1, ERR)) win = 0;      // :609
            else if (maxsum1 > __fadd_rn(maxsum0, ERR)) win = 1; // :625
            if (win >= 0) {
                const float target = win == 0 ? maxsum0 : maxsum1;
                Uniq uB; uB.init();
                M.reset();
                while (M.next(R, g, s, pres, bad)) {
                    if (g >= P.n_refs) continue;
                    bool inI = false;
#pragma unroll
                    for (int i = 0; i < NF; i++)
                        if (((top & pres) >> i) & 1u)
                            if (__fsub_rn(R.mx[i], s[i]) < ERR) inI = true;
                    if (!inI) continue;
                    float p0, p1;
                    pair_sums<NF>(s, p0, p1);
                    if ((win == 0 ? p0 : p1) == target) uB.add(P.v[g]);   // :611-612, :627-628
                }
                if (uB.single()) { type = 'C'; tax = uB.val; sim = target; done = true; }
            }
        }
        // ---- pass 3 : stage C (Exam_2) ----
        if (!done) {
            float h; int k;                                      // k: 0, 1, or 2 = tie (:230-253)
            if (hi0 > hi1) { h = hi0; k = 0; } else if (hi0 < hi1) { h = hi1; k = 1; } else { h = hi0; k = 2; }
            Uniq uV; uV.init();
            Uniq uM[LIME_RANKS];
#pragma unroll
            for (int x = 0; x < LIME_RANKS; x++) uM[x].init();
            if (__fsub_rn(h, 0.0f) < ERR) {
                // every genome (present or absent) is within ERROR of h
                uV.any = true; uV.same = P.all_v_same != 0; uV.val = P.all_v_tax;
#pragma unroll
                for (int x = 0; x < LIME_RANKS; x++) {
                    uM[x].any = true; uM[x].same = P.all_m_same[x] != 0; uM[x].val = P.all_m_tax[x];
                }
            } else {
                M.reset();
                while (M.next(R, g, s, pres, bad)) {
                    if (g >= P.n_refs) continue;
                    float p0, p1;
                    pair_sums<NF>(s, p0, p1);
                    const bool c0 = __fsub_rn(h, p0) < ERR, c1 = __fsub_rn(h, p1) < ERR;
                    const bool in = k == 0 ? c0 : k == 1 ? c1 : (c0 || c1);
                    if (!in) continue;
                    uV.add(P.v[g]);
                    if (P.higher && P.m) {
#pragma unroll
                        for (int x = 0; x < LIME_RANKS; x++)
                            if (x >= P.tax_rank - 1) uM[x].add(P.m[(size_t)x * P.n_refs + g]);
                    }
                }
            }
            if (uV.single()) { type = 'C'; tax = uV.val; sim = h; }        // :652-670
            else {
                type = 'A';                                                // :679-684
                if (P.higher && P.m) {                                     // :262-302
                    bool found = false;
#pragma unroll
                    for (int x = 0; x < LIME_RANKS; x++) {
                        if (found || x < P.tax_rank - 1) continue;
                        if (uM[x].single() && uM[x].val != 0) { type = 'H'; tax = uM[x].val; sim = h; found = true; }
                    }
                }
            }
        }
    }
    P.out_type[r] = type; P.out_tax[r] = tax; P.out_sim[r] = sim;
    if (bad) atomicOr(P.err, 1u);
}

} // namespace cls

cudaError_t launch_classify(const ClsArgs &a, cudaStream_t s)
{
    if (a.n_reads == 0) return cudaSuccess;
    const unsigned grid = (a.n_reads + 127) / 128;
    if (a.n_files == 4) cls::classify_kernel<4><<<grid, 128, 0, s>>>(a);
    else                cls::classify_kernel<2><<<grid, 128, 0, s>>>(a);
    return cudaGetLastError();
}

} // namespace lime
