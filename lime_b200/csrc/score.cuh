// score.cuh -- scoring of one (read, genome) pair inside one cluster and the accumulator update,
// shared by the fused kernel (k12_stream.cu) and the cluster-list kernels (k2_similarity.cu).
#pragma once
#include "common.cuh"

namespace lime {

// symbol -> class, ClusterBWT_DA.cpp:455-470; every unlisted byte is class 0 because
// unordered_map::operator[] default-inserts 0 (:96,123)
struct ClassTable { uint8_t v[256]; };
constexpr ClassTable make_class_table()
{
    ClassTable t{};
    const char sym[16] = "ACGTRYSWKMBDHVN";
    for (int c = 0; c < 15; c++) t.v[(unsigned char)sym[c]] = (uint8_t)c;
    t.v[0] = 15;
    return t;
}
static __device__ const ClassTable d_class_table = make_class_table();

__device__ __forceinline__ uint32_t class_of(uint8_t sym) { return __ldg(&d_class_table.v[sym]); }

// count vectors are 16 uint8 packed little-endian into a uint4 (class c = byte c&3 of word c>>2)

// no IUPAC class (4..14) present on either side: t = sum_c min(R[c], G[c])  (:133-144)
__device__ __forceinline__ uint32_t score_plain(const uint4 &R, const uint4 &G)
{
    return __vsadu4(__vminu4(R.x, G.x), 0u) + __vsadu4(__vminu4(R.y, G.y), 0u) +
           __vsadu4(__vminu4(R.z, G.z), 0u) + __vsadu4(__vminu4(R.w, G.w), 0u);
}
// only classes 0..3 present anywhere in the cluster
__device__ __forceinline__ uint32_t score_acgt(uint32_t Rx, uint32_t Gx)
{
    return __vsadu4(__vminu4(Rx, Gx), 0u);
}

__device__ __forceinline__ uint32_t byte_of(const uint4 &v, int c)
{
    const uint32_t w = (c >> 2) == 0 ? v.x : (c >> 2) == 1 ? v.y : (c >> 2) == 2 ? v.z : v.w;
    return (w >> ((c & 3) * 8)) & 0xffu;
}

// class a (4..14) compatible with base i (0..3), ClusterBWT_DA.cpp:472-487, as bit i of entry a
__device__ __forceinline__ constexpr unsigned compat_bits(int a)
{
    return a == 4 ? 0x5u : a == 5 ? 0xau : a == 6 ? 0x6u : a == 7 ? 0x9u : a == 8 ? 0xcu :
           a == 9 ? 0x3u : a == 10 ? 0xeu : a == 11 ? 0xdu : a == 12 ? 0xbu : a == 13 ? 0x7u :
           a == 14 ? 0xfu : 0u;
}

// full matcher incl. the IUPAC cross loop (:146-177).  In the "genome IUPAC vs read base" branch
// the reference zeroes one operand before subtracting it from the other (:153-154, :158-159), so
// the other operand keeps its value; t wraps as uint8 (:131), taken mod 256 by the caller.
static __device__ __noinline__ uint32_t score_iupac(const uint4 R, const uint4 G)
{
    uint32_t rr[16], gg[16], t = 0;
#pragma unroll
    for (int c = 0; c < 16; c++) {
        const uint32_t r = byte_of(R, c), g = byte_of(G, c);
        if (r > g) { t += g; rr[c] = r - g; gg[c] = 0; }
        else       { t += r; gg[c] = g - r; rr[c] = 0; }
    }
#pragma unroll
    for (int i = 0; i < 4; i++) {
#pragma unroll
        for (int a = 4; a < 15; a++) {
            if (!((compat_bits(a) >> i) & 1u)) continue;
            if (gg[a] > 0) {
                if (gg[a] > rr[i]) { t += rr[i]; rr[i] = 0; }
                else               { t += gg[a]; gg[a] = 0; }
            }
            if (rr[a] > 0) {
                if (rr[a] > gg[i]) { t += gg[i]; rr[a] -= gg[i]; gg[i] = 0; }
                else               { t += rr[a]; gg[i] -= rr[a]; rr[a] = 0; }
            }
        }
    }
    return t;
}

// ---- accumulator update ----------------------------------------------------------------------
// Dense [n_reads][pitch] uint8 cells updated with a 32-bit atomic on the containing word; an add
// that carries out of its byte raises *carry and the host reruns the pass into exact 32-bit
// cells for a band of reads [lo, hi) (reduced mod 256 afterwards = the uint8 wrap of
// SimArray_, ClusterBWT_DA.cpp:183,247).
struct SimAcc {
    uint8_t *sim8; uint64_t pitch8;
    uint32_t *sim32; uint32_t lo, hi, n_refs;
    unsigned int *carry;
};

template <bool EXACT32>
__device__ __forceinline__ void sim_add(const SimAcc &P, uint32_t r, uint32_t g, uint32_t t)
{
    if (EXACT32) {
        if (r >= P.lo && r < P.hi) atomicAdd(&P.sim32[(size_t)(r - P.lo) * P.n_refs + g], t);
    } else {
        uint8_t *row = P.sim8 + (size_t)r * P.pitch8;
        unsigned *w = reinterpret_cast<unsigned *>(row) + (g >> 2);
        const unsigned sh = (g & 3u) * 8u;
        const unsigned old = atomicAdd(w, t << sh);
        if (((old >> sh) & 0xffu) + t > 0xffu) *P.carry = 1u;
    }
}

} // namespace lime
