// host_narrow.cpp -- host half of the narrowed lcp transfer of lime_load(LIME_MEM_HOST).
//
// Every step of the path reads lcp only through `lcp[i] < alpha` (ClusterLCP.cpp:150-202, :246-264),
// so for alpha <= 255 the byte min(lcp, 255) decides every test exactly as the 32-bit value does.
// Shipping that byte takes 1 instead of 4 of the 9 bytes per position over PCIe.  No CUDA in here:
// plain C++ (g++, AVX2 clone picked at load time), tested on the CPU (tests/test_narrow_lcp.py).
#include <algorithm>
#include <atomic>
#include <immintrin.h>
#include <cstdint>
#include <cstdlib>
#include <chrono>
#include <deque>
#include <mutex>
#include <thread>
#include <sched.h>
#include <vector>
#include "../../include/lime_cuda.h"
#include "narrow_loop.h"

namespace {

__attribute__((target_clones("avx2", "default")))
void narrow_range(const uint32_t *__restrict__ src, uint8_t *__restrict__ dst, uint64_t n)
{
    for (uint64_t i = 0; i < n; i++) {
        const uint32_t v = src[i];
        dst[i] = (uint8_t)(v < 255u ? v : 255u);
    }
}

// bit j of dst[i] = (src[8 i + j] < alpha): the only thing the path ever asks of lcp, once alpha is known.
// n positions, n a multiple of 8 except for the last call of an array; positions beyond n read as heads (1).
void pack_range_scalar(const uint32_t *src, uint8_t *dst, uint64_t n, uint32_t alpha)
{
    for (uint64_t i = 0; i < n; i += 8) {
        unsigned b = 0;
        for (unsigned j = 0; j < 8; j++) b |= (unsigned)(i + j >= n || src[i + j] < alpha) << j;
        dst[i >> 3] = (uint8_t)b;
    }
}
__attribute__((target("avx2")))
void pack_range_avx2(const uint32_t *src, uint8_t *dst, uint64_t n, uint32_t alpha)
{
    const __m256i flip = _mm256_set1_epi32((int)0x80000000u);
    const __m256i a = _mm256_xor_si256(_mm256_set1_epi32((int)alpha), flip);
    uint64_t i = 0;
    for (; i + 32 <= n; i += 32) {                    // 4 bytes of bits per round
        uint32_t w = 0;
        for (int k = 0; k < 4; k++) {
            const __m256i v = _mm256_xor_si256(_mm256_loadu_si256(reinterpret_cast<const __m256i *>(src + i + 8 * k)), flip);
            w |= (uint32_t)_mm256_movemask_ps(_mm256_castsi256_ps(_mm256_cmpgt_epi32(a, v))) << (8 * k);
        }
        dst[(i >> 3) + 0] = (uint8_t)w; dst[(i >> 3) + 1] = (uint8_t)(w >> 8);
        dst[(i >> 3) + 2] = (uint8_t)(w >> 16); dst[(i >> 3) + 3] = (uint8_t)(w >> 24);
    }
    if (i < n) pack_range_scalar(src + i, dst + (i >> 3), n - i, alpha);
}
void pack_range(const uint32_t *src, uint8_t *dst, uint64_t n, uint32_t alpha)
{
    static const bool avx2 = __builtin_cpu_supports("avx2");
    if (avx2) pack_range_avx2(src, dst, n, alpha); else pack_range_scalar(src, dst, n, alpha);
}

// dst[3 i .. 3 i + 2] = the low three bytes of src[i] (document ids below 2^24), little endian
void pack3_scalar(const uint32_t *src, uint8_t *dst, uint64_t n)
{
    for (uint64_t i = 0; i < n; i++) {
        const uint32_t v = src[i];
        dst[3 * i] = (uint8_t)v; dst[3 * i + 1] = (uint8_t)(v >> 8); dst[3 * i + 2] = (uint8_t)(v >> 16);
    }
}
__attribute__((target("avx2")))
void pack3_avx2(const uint32_t *src, uint8_t *dst, uint64_t n)
{
    uint64_t i = 0;
    if ((reinterpret_cast<uintptr_t>(dst) & 15) == 0) {
        // 16 positions -> 48 bytes = three aligned 16-byte non-temporal stores: the stage buffer is written once and
        // read by the DMA engine, so the lines need not be fetched into the cache first
        const __m128i sh = _mm_setr_epi8(0, 1, 2, 4, 5, 6, 8, 9, 10, 12, 13, 14, -1, -1, -1, -1);
        for (; i + 16 <= n; i += 16) {
            const __m128i a = _mm_shuffle_epi8(_mm_loadu_si128(reinterpret_cast<const __m128i *>(src + i)), sh);
            const __m128i b = _mm_shuffle_epi8(_mm_loadu_si128(reinterpret_cast<const __m128i *>(src + i + 4)), sh);
            const __m128i c = _mm_shuffle_epi8(_mm_loadu_si128(reinterpret_cast<const __m128i *>(src + i + 8)), sh);
            const __m128i d = _mm_shuffle_epi8(_mm_loadu_si128(reinterpret_cast<const __m128i *>(src + i + 12)), sh);
            __m128i *o = reinterpret_cast<__m128i *>(dst + 3 * i);
            _mm_stream_si128(o + 0, _mm_or_si128(a, _mm_slli_si128(b, 12)));                       // a[0..12) b[0..4)
            _mm_stream_si128(o + 1, _mm_or_si128(_mm_srli_si128(b, 4), _mm_slli_si128(c, 8)));     // b[4..12) c[0..8)
            _mm_stream_si128(o + 2, _mm_or_si128(_mm_srli_si128(c, 8), _mm_slli_si128(d, 4)));     // c[8..12) d[0..12)
        }
        _mm_sfence();           // non-temporal stores are ordered before the release store that publishes the slice
    } else {
        const __m256i sh = _mm256_setr_epi8(0, 1, 2, 4, 5, 6, 8, 9, 10, 12, 13, 14, -1, -1, -1, -1,
                                            0, 1, 2, 4, 5, 6, 8, 9, 10, 12, 13, 14, -1, -1, -1, -1);
        // a round stores 2 x 16 bytes for 24: the 4 bytes of slack fall on positions this call writes afterwards, so
        // the last 8 .. 15 positions are left to the scalar loop and nothing is written beyond 3 n
        for (; i + 16 <= n; i += 8) {
            const __m256i v = _mm256_shuffle_epi8(_mm256_loadu_si256(reinterpret_cast<const __m256i *>(src + i)), sh);
            _mm_storeu_si128(reinterpret_cast<__m128i *>(dst + 3 * i), _mm256_castsi256_si128(v));
            _mm_storeu_si128(reinterpret_cast<__m128i *>(dst + 3 * i + 12), _mm256_extracti128_si256(v, 1));
        }
    }
    pack3_scalar(src + i, dst + 3 * i, n - i);
}
void pack3(const uint32_t *src, uint8_t *dst, uint64_t n)
{
    static const bool avx2 = __builtin_cpu_supports("avx2");
    if (avx2) pack3_avx2(src, dst, n); else pack3_scalar(src, dst, n);
}

} // namespace

extern "C" int lime_host_threads(void)
{
    if (const char *e = getenv("LIME_HOST_THREADS")) {
        const int t = atoi(e);
        if (t >= 1) return std::min(t, 256);
    }
    // the cores this process may run on (a container's cpuset), not the machine's
    unsigned hc = std::thread::hardware_concurrency();
    cpu_set_t set;
    if (sched_getaffinity(0, sizeof(set), &set) == 0 && CPU_COUNT(&set) > 0) hc = (unsigned)CPU_COUNT(&set);
    return (int)std::max(1u, std::min(hc ? hc : 1u, 16u));
}

extern "C" int lime_host_narrow_lcp(const uint32_t *lcp, uint8_t *out, uint64_t n, int threads)
{
    if ((!lcp || !out) && n) return LIME_ERR_ARG;
    if (threads <= 0) threads = lime_host_threads();
    // below a few MB one thread is faster than starting the others
    const uint64_t per = std::max<uint64_t>((n + threads - 1) / threads, 1u << 20);
    std::vector<std::thread> pool;
    for (uint64_t a = per; a < n; a += per)
        pool.emplace_back(narrow_range, lcp + a, out + a, std::min(per, n - a));
    narrow_range(lcp, out, std::min(per, n));
    for (auto &t : pool) t.join();
    return LIME_OK;
}

extern "C" int lime_host_pack_lcp(const uint32_t *lcp, uint8_t *out, uint64_t n, uint32_t alpha, int threads)
{
    if ((!lcp || !out) && n) return LIME_ERR_ARG;
    if (threads <= 0) threads = lime_host_threads();
    const uint64_t per = std::max<uint64_t>(((n + threads - 1) / threads + 511) / 512 * 512, 1u << 20);
    std::vector<std::thread> pool;
    for (uint64_t a = per; a < n; a += per)
        pool.emplace_back(pack_range, lcp + a, out + (a >> 3), std::min(per, n - a), alpha);
    pack_range(lcp, out, std::min(per, n), alpha);
    for (auto &t : pool) t.join();
    return LIME_OK;
}

// ---------------------------------------------------------------------------------------------
// The chunk pipeline of lime_load: `threads` workers narrow chunk after chunk into a ring of
// `slots` page-locked stage buffers; the calling thread (lime_abi_load.cu) takes each finished chunk,
// queues its copy and gives the slot back once the copy's event has completed.  Workers only touch
// host memory; all CUDA calls stay on the calling thread.
// ---------------------------------------------------------------------------------------------
namespace lime {

struct NarrowPipe {
    const uint32_t *src = nullptr;
    uint64_t n = 0, chunk = 0, n_chunks = 0;
    int threads = 0, slots = 0;
    uint8_t *const *stage = nullptr;                 // [slots] buffers of `chunk` bytes (bits: chunk / 8)
    uint32_t bits_alpha = 0;                         // 0: bytes min(lcp, 255); else one bit lcp < bits_alpha
    const uint32_t *da3 = nullptr;                   // non-null: da of the chunk as 3 bytes per position at stage + chunk
    std::vector<std::atomic<int>> done;              // per chunk: workers that finished their slice
    std::atomic<uint64_t> free_upto{0};              // chunks < free_upto own a free slot
    std::vector<std::thread> pool;

    NarrowPipe(const uint32_t *src_, uint64_t n_, uint64_t chunk_, int threads_, int slots_, uint8_t *const *stage_, uint32_t bits_alpha_,
               const uint32_t *da3_)
        : src(src_), n(n_), chunk(chunk_), n_chunks((n_ + chunk_ - 1) / chunk_), threads(threads_), slots(slots_),
          stage(stage_), bits_alpha(bits_alpha_), da3(da3_), done((n_ + chunk_ - 1) / chunk_)
    {
        for (auto &d : done) d.store(0, std::memory_order_relaxed);
        free_upto.store((uint64_t)slots, std::memory_order_release);
        for (int t = 0; t < threads; t++) pool.emplace_back([this, t] { work(t); });
    }
    ~NarrowPipe()
    {
        free_upto.store(~0ull, std::memory_order_release);      // an aborted load must not strand the workers
        for (auto &t : pool) t.join();
    }
    void work(int t)
    {
        for (uint64_t c = 0; c < n_chunks; c++) {
            while (c >= free_upto.load(std::memory_order_acquire)) std::this_thread::yield();
            const uint64_t base = c * chunk, len = std::min(chunk, n - base);
            // slices on 64-byte boundaries of the output (bits: 512 positions)
            const uint64_t unit = bits_alpha ? 512 : 64;
            const uint64_t per = ((len + threads - 1) / threads + unit - 1) / unit * unit;
            const uint64_t a = std::min(len, per * (uint64_t)t), b = std::min(len, a + per);
            if (b > a) {
                if (bits_alpha) pack_range(src + base + a, stage[c % slots] + (a >> 3), b - a, bits_alpha);
                else narrow_range(src + base + a, stage[c % slots] + a, b - a);
                if (da3) pack3(da3 + base + a, stage[c % slots] + chunk + 3 * a, b - a);
            }
            done[c].fetch_add(1, std::memory_order_release);
        }
    }
    bool ready(uint64_t c) const { return done[c].load(std::memory_order_acquire) == threads; }
    // the copies of chunks < k have completed: chunks < k + slots may be narrowed
    void released(uint64_t k) { free_upto.store(k + (uint64_t)slots, std::memory_order_release); }
};

NarrowPipe *narrow_pipe_open(const uint32_t *src, uint64_t n, uint64_t chunk, int threads, int slots, uint8_t *const *stage,
                             uint32_t bits_alpha, const uint32_t *da3)
{
    return new NarrowPipe(src, n, chunk, threads, slots, stage, bits_alpha, da3);
}
bool narrow_pipe_ready(const NarrowPipe *p, uint64_t c) { return p->ready(c); }
void narrow_pipe_released(NarrowPipe *p, uint64_t k) { p->released(k); }
void narrow_pipe_close(NarrowPipe *p) { delete p; }

} // namespace lime

// CPU test hook: narrow_transfer (narrow_loop.h, the loop lime_load runs) over a simulated copy queue -- an
// in-order worker that performs each queued copy `copy_delay_us` after taking it, so a stage slot that
// is rewritten before its copy has run shows up as wrong output.
namespace {

struct SimQueue {
    struct Job { int slot; uint64_t base, len, seq; };
    uint8_t *out; uint8_t *const *stage; int delay_us; bool bits;
    uint8_t *out_da3 = nullptr; uint64_t chunk = 0;
    std::mutex mu; std::deque<Job> jobs; bool stop = false;
    std::atomic<uint64_t> recorded[8], completed[8];
    std::thread worker;
    SimQueue(uint8_t *out_, uint8_t *const *stage_, int delay_us_, bool bits_) : out(out_), stage(stage_), delay_us(delay_us_), bits(bits_)
    {
        for (int s = 0; s < 8; s++) { recorded[s].store(0); completed[s].store(0); }
        worker = std::thread([this] { run(); });
    }
    ~SimQueue()
    {
        { std::lock_guard<std::mutex> g(mu); stop = true; }
        worker.join();
    }
    void run()
    {
        for (;;) {
            Job j;
            {
                std::lock_guard<std::mutex> g(mu);
                if (jobs.empty()) { if (stop) return; j.slot = -1; }
                else { j = jobs.front(); jobs.pop_front(); }
            }
            if (j.slot < 0) { std::this_thread::yield(); continue; }
            if (delay_us > 0) std::this_thread::sleep_for(std::chrono::microseconds(delay_us));
            if (bits) std::copy(stage[j.slot], stage[j.slot] + (j.len + 7) / 8, out + j.base / 8);
            else std::copy(stage[j.slot], stage[j.slot] + j.len, out + j.base);
            if (out_da3) std::copy(stage[j.slot] + chunk, stage[j.slot] + chunk + 3 * j.len, out_da3 + 3 * j.base);
            completed[j.slot].store(j.seq, std::memory_order_release);
        }
    }
    int stage_copy(int slot, uint64_t base, uint64_t len)
    {
        const uint64_t seq = recorded[slot].load(std::memory_order_relaxed) + 1;
        recorded[slot].store(seq, std::memory_order_relaxed);
        std::lock_guard<std::mutex> g(mu);
        jobs.push_back({slot, base, len, seq});
        return 0;
    }
    int slot_state(int slot)
    {
        return completed[slot].load(std::memory_order_acquire) == recorded[slot].load(std::memory_order_relaxed) ? 0 : 1;
    }
    int rest(uint64_t, uint64_t) { return 0; }
    void drain()
    {
        for (;;) {
            { std::lock_guard<std::mutex> g(mu); if (jobs.empty()) break; }
            std::this_thread::yield();
        }
    }
};

} // namespace

static int pipelined(const uint32_t *lcp, uint8_t *out, uint64_t n, uint64_t chunk, int threads, int slots, int copy_delay_us,
                     uint32_t bits_alpha, const uint32_t *da = nullptr, uint8_t *out_da3 = nullptr)
{
    if ((da != nullptr) != (out_da3 != nullptr)) return LIME_ERR_ARG;
    if ((!lcp || !out) && n) return LIME_ERR_ARG;
    if (chunk == 0 || slots < 1 || slots > 8 || (bits_alpha && chunk % 512)) return LIME_ERR_ARG;
    if (threads <= 0) threads = lime_host_threads();
    if (n == 0) return LIME_OK;
    std::vector<std::vector<uint8_t>> bufs(slots, std::vector<uint8_t>(da ? 4 * chunk : chunk));
    std::vector<uint8_t *> stage;
    for (auto &b : bufs) stage.push_back(b.data());
    int rc;
    {
        SimQueue q(out, stage.data(), copy_delay_us, bits_alpha != 0);
        q.out_da3 = out_da3; q.chunk = chunk;
        rc = lime::narrow_transfer(q, lcp, n, chunk, threads, slots, stage.data(), bits_alpha, da);
        q.drain();
    }                                                      // joins the worker: every queued copy has run
    return rc ? LIME_ERR_CUDA : LIME_OK;
}

extern "C" int lime_host_narrow_lcp_pipelined(const uint32_t *lcp, uint8_t *out, uint64_t n, uint64_t chunk,
                                              int threads, int slots, int copy_delay_us)
{
    return pipelined(lcp, out, n, chunk, threads, slots, copy_delay_us, 0);
}

extern "C" int lime_host_pack_lcp_pipelined(const uint32_t *lcp, uint8_t *out, uint64_t n, uint64_t chunk, uint32_t alpha,
                                            int threads, int slots, int copy_delay_us)
{
    if (alpha == 0) return LIME_ERR_ARG;
    return pipelined(lcp, out, n, chunk, threads, slots, copy_delay_us, alpha);
}

extern "C" int lime_host_pack_da3(const uint32_t *da, uint8_t *out, uint64_t n)
{
    if ((!da || !out) && n) return LIME_ERR_ARG;
    pack3(da, out, n);
    return LIME_OK;
}

extern "C" int lime_host_pack_lcp_da3_pipelined(const uint32_t *lcp, const uint32_t *da, uint8_t *out_bits, uint8_t *out_da3, uint64_t n,
                                                uint64_t chunk, uint32_t alpha, int threads, int slots, int copy_delay_us)
{
    if (alpha == 0 || ((!da || !out_da3) && n)) return LIME_ERR_ARG;
    return pipelined(lcp, out_bits, n, chunk, threads, slots, copy_delay_us, alpha, da, out_da3);
}
