// narrow_loop.h -- the chunk loop of the narrowed lcp transfer, independent of CUDA so that the CPU tests
// run the very loop lime_load runs (host_narrow.cpp drives it with a simulated asynchronous copy queue,
// lime_abi_load.cu with cudaMemcpyAsync + events).
#pragma once
#include <algorithm>
#include <cstdint>
#include <thread>

namespace lime {

struct NarrowPipe;
// bits_alpha = 0: a chunk is staged as the bytes min(lcp, 255); else as one bit per position, lcp < bits_alpha
// (bit j of byte i = position 8 i + j of the chunk; chunk must be a multiple of 512)
// da3 non-null: the workers also stage da of the chunk as 3 bytes per position (ids below 2^24) at stage + chunk
// (a stage buffer then holds 4 * chunk bytes)
NarrowPipe *narrow_pipe_open(const uint32_t *src, uint64_t n, uint64_t chunk, int threads, int slots, uint8_t *const *stage,
                             uint32_t bits_alpha, const uint32_t *da3);
bool narrow_pipe_ready(const NarrowPipe *p, uint64_t c);
void narrow_pipe_released(NarrowPipe *p, uint64_t k);
void narrow_pipe_close(NarrowPipe *p);

// Q is the copy queue (in-order, asynchronous):
//   int stage_copy(int slot, uint64_t base, uint64_t len)  queue "stage[slot][0, len) -> position base" and
//                                                          mark the slot pending until that copy is done
//   int slot_state(int slot)                               0 = its last stage_copy is done, 1 = pending, < 0 = error
//   int rest(uint64_t base, uint64_t len)                  queue whatever else belongs to the chunk
// Returns 0 or the first error of Q.  On return every chunk has been queued; the last `slots` copies may
// still be pending (the stage buffers must outlive them).
template <class Q>
int narrow_transfer(Q &q, const uint32_t *lcp, uint64_t n, uint64_t chunk, int threads, int slots, uint8_t *const *stage,
                    uint32_t bits_alpha = 0, const uint32_t *da3 = nullptr)
{
    NarrowPipe *pipe = narrow_pipe_open(lcp, n, chunk, threads, slots, stage, bits_alpha, da3);
    struct Closer { NarrowPipe *p; ~Closer() { narrow_pipe_close(p); } } closer{pipe};
    const uint64_t n_chunks = (n + chunk - 1) / chunk;
    uint64_t queued = 0, released = 0;       // chunks whose copy is queued / known to be done
    int bad = 0;
    auto poll = [&]() {
        while (released < queued) {
            const int st = q.slot_state((int)(released % (uint64_t)slots));
            if (st == 1) break;
            if (st < 0) { bad = st; break; }
            released++;
        }
        narrow_pipe_released(pipe, released);   // chunks < released + slots may be narrowed
    };
    for (uint64_t c = 0; c < n_chunks; c++) {
        while (!narrow_pipe_ready(pipe, c)) {
            poll();
            if (bad) return bad;
            std::this_thread::yield();
        }
        const uint64_t base = c * chunk, len = std::min(chunk, n - base);
        int rc = q.stage_copy((int)(c % (uint64_t)slots), base, len);
        if (rc) return rc;
        queued = c + 1;
        rc = q.rest(base, len);
        if (rc) return rc;
        poll();
        if (bad) return bad;
    }
    return 0;
}

} // namespace lime
