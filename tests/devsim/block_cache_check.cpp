// Test tooling: librtb's block cache (csrc/rtb_block_cache.h) on the host (emul/cuda_emulation.h counts the frees per device).
// What must hold: a released block is handed to the next request of about its size on the same device; at most 8 blocks per
// device and RTB_BLOCK_CACHE_MB in total; a thread that deferred its frees (a rank thread of rtb_render_multi) never calls
// cudaFree -- what it drops is freed by flush_deferred(); without deferral a drop only frees on the block's own device.
#define RTB_BUILD_EMULATE 1
#include "../../raytracing-hw_b200/csrc/rtb_block_cache.h"

#include <cstdio>
#include <thread>

static int frees_total() { int t = 0; for (int f : emu_frees) t += f; return t; }

int main() {
    setenv("RTB_BLOCK_CACHE_MB", "64", 1);
    rtb::BlockCache c;
    int bad = 0;
    auto expect = [&](bool ok, const char *what) { if (!ok) { std::printf("FAILED: %s\n", what); bad = 1; } };
    void *p = nullptr, *q = nullptr;
    size_t got = 0;
    // reuse on the same device, not on another one, not for a much smaller request
    cudaSetDevice(0);
    expect(c.acquire(0, 10000000, &p, &got) == cudaSuccess && got == 10000000, "fresh block");
    c.release(0, p, got);
    expect(c.acquire(1, 9000000, &q, &got) == cudaSuccess && q != p, "another device does not get the block");
    c.release(1, q, got);
    expect(c.acquire(0, 1000000, &q, &got) == cudaSuccess && q != p && got == 1000000, "a much smaller request does not take the block");
    c.release(0, q, got);
    expect(c.acquire(0, 9000000, &q, &got) == cudaSuccess && q == p && got == 10000000, "same device, about the size: reused");
    c.release(0, q, got);
    expect(frees_total() == 0, "nothing freed so far");
    // at most 8 blocks per device: the oldest of THAT device goes, on that device
    for (int i = 0; i < 12; ++i) { c.acquire(2, 5000 + i, &p, &got); c.release(2, p, got); }
    size_t on2 = 0;
    for (auto &b : c.blocks) on2 += b.device == 2;
    expect(on2 <= 8 && emu_frees[2] >= 1 && frees_total() == emu_frees[2], "per-device block limit, freed on the block's device");
    // a deferring thread never frees
    const int before = frees_total();
    std::thread t([&] {
        rtb::BlockCache::defer_flag() = true;
        for (int i = 0; i < 20; ++i) {
            void *a = nullptr;
            size_t g = 0;
            c.acquire(3, 7000 + i, &a, &g);   // (a fresh allocation: nothing cached for device 3)
            c.release(3, a, g);
        }
        // over the byte limit: blocks of OTHER devices would have to go -- deferred as well
        void *big = nullptr;
        size_t g = 0;
        c.acquire(3, (size_t) 60 << 20, &big, &g);
        c.release(3, big, g);
    });
    t.join();
    expect(frees_total() == before, "a thread that defers its frees calls cudaFree for nothing");
    expect(!c.deferred.empty(), "its drops wait in the deferred list");
    c.flush_deferred();
    expect(c.deferred.empty() && frees_total() > before, "flush_deferred frees them");
    expect(c.total <= c.limit(), "the byte limit holds");
    // out of memory: a plain thread gives the cached blocks of the device back and tries again; a deferring thread does
    // neither (trim is cudaFree) and reports the error -- rtb_render_multi trims after its threads have joined
    c.acquire(0, 3000000, &p, &got);
    c.release(0, p, got);   // something parked on device 0
    {
        const int f0 = frees_total();
        emu_malloc_fail_next = 1;
        expect(c.acquire(0, 20000000, &q, &got) == cudaSuccess && frees_total() > f0, "out of memory: the cache is trimmed and the request retried");
        c.release(0, q, got);
        const int f1 = frees_total();
        cudaError_t e_thread = cudaSuccess;
        std::thread t2([&] {
            rtb::BlockCache::defer_flag() = true;
            void *a = nullptr;
            size_t g = 0;
            emu_malloc_fail_next = 1;
            e_thread = c.acquire(0, 50000000, &a, &g);
        });
        t2.join();
        emu_malloc_fail_next = 0;
        expect(e_thread != cudaSuccess && frees_total() == f1, "out of memory on a deferring thread: an error, no cudaFree");
    }
    // an oversized block is never kept
    c.acquire(0, (size_t) 80 << 20, &p, &got);
    const int f0 = frees_total();
    c.release(0, p, got);
    expect(frees_total() == f0 + 1, "a block above the byte limit goes straight back");
    c.trim(-1);
    expect(c.blocks.empty() && c.total == 0, "trim empties the cache");
    std::printf("%s\n", bad ? "FAILED" : "ok");
    return bad;
}
