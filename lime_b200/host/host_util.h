// host_util.h -- small helpers shared by the drop-in executables (ClusterLCP, ClusterBWT_DA,
// Classify, LiME_fused).  They keep the reference's file naming and error behaviour:
// message on stderr, "fopen failed, errno = N" on stdout, exit(EXIT_FAILURE).
#pragma once
#include <cerrno>
#include <chrono>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <iostream>
#include <string>
#include "../../include/lime_cuda.h"
#include "../../include/lime_host.h"

namespace limehost {

// <stem> = fileFasta truncated at the first ".fasta" (ClusterLCP.cpp:294, ClusterBWT_DA.cpp:536)
inline std::string stem_of(const std::string &fasta) { return fasta.substr(0, fasta.find(".fasta")); }

[[noreturn]] inline void die_open(const std::string &path)
{
    std::cerr << "Error opening " << path << "." << std::endl;
    printf("fopen failed, errno = %d\n", errno);
    exit(EXIT_FAILURE);
}

[[noreturn]] inline void die_lime(lime_ctx *ctx, const char *what, int rc)
{
    std::cerr << "Error: " << what << " failed (" << rc << "): " << lime_last_error(ctx) << std::endl;
    exit(EXIT_FAILURE);
}

inline uint64_t file_size(const std::string &path)
{
    FILE *f = fopen(path.c_str(), "rb");
    if (!f) die_open(path);
    fseek(f, 0, SEEK_END);
    const uint64_t n = (uint64_t)ftell(f);
    fclose(f);
    return n;
}

// whole file into page-locked memory (full-rate, asynchronous H2D); caller frees with lime_pinned_free
inline void *read_file_pinned(const std::string &path, uint64_t *bytes)
{
    FILE *f = fopen(path.c_str(), "rb");
    if (!f) die_open(path);
    fseek(f, 0, SEEK_END);
    const uint64_t n = (uint64_t)ftell(f);
    fseek(f, 0, SEEK_SET);
    void *p = lime_pinned_alloc(n ? n : 1);
    if (!p) { std::cerr << "Error: cannot allocate " << n << " bytes of pinned memory" << std::endl; exit(EXIT_FAILURE); }
    uint64_t got = 0;
    while (got < n) {
        const size_t r = fread(static_cast<char *>(p) + got, 1, (size_t)std::min<uint64_t>(n - got, 1ull << 30), f);
        if (r == 0) break;
        got += r;
    }
    fclose(f);
    if (got != n) { std::cerr << "Error reading " << path << "." << std::endl; exit(EXIT_FAILURE); }
    *bytes = n;
    return p;
}

struct Timer {
    std::chrono::steady_clock::time_point t0 = std::chrono::steady_clock::now();
    double seconds() const { return std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count(); }
};

// the `threads` argument of the reference CLIs keeps its position; one process drives one GPU, so it
// only selects the device ordinal through LIME_DEVICE (default 0) and is otherwise reported back
inline int device_from_env()
{
    const char *e = getenv("LIME_DEVICE");
    return e ? atoi(e) : 0;
}

// The executables know alpha from their command line before they read a file: they name it to the library
// (lime_set_alpha), so that lime_load ships lcp as the bit lcp < alpha when alpha <= 255 -- 5.1 instead of
// 9 bytes per position over PCIe, identical results.  LIME_NARROW_LCP=0 switches it off, LIME_NARROW_LCP=8 ships
// the bytes min(lcp, 255) instead (lime_set_alpha_limit).
inline void alpha_limit_from_env(lime_ctx *ctx, uint32_t alpha)
{
    const char *e = getenv("LIME_NARROW_LCP");
    if (e && atoi(e) == 8) lime_set_alpha_limit(ctx, alpha);
    else if (!e || atoi(e) > 0) lime_set_alpha(ctx, alpha);
}

} // namespace limehost
