// The block cache of librtb's device allocations (csrc/rtb_api.cu), as a header so that tests/devsim/block_cache_check.cpp can
// drive it on the host (emul/cuda_emulation.h).
#pragma once
#ifdef RTB_BUILD_EMULATE
#include "cuda_emulation.h"
#else
#include <cuda_runtime.h>
#endif

#include <algorithm>
#include <cstdlib>
#include <mutex>
#include <vector>

// Device blocks of destroyed scenes (scene arena, path state, staging) are kept -- a few, bounded -- and handed to the
// next scene that asks for a block of about that size.  A frame server that uploads a scene per frame otherwise pays
// cudaMalloc / cudaFree of ~1 GB per frame: cudaFree alone was seen at 0.2 - 1.1 s (profiles/r01m_e2e_probe.txt).
// RTB_BLOCK_CACHE_MB: the most the cache keeps (default 8192; 0: every block goes straight back to the driver).
namespace rtb {
struct BlockCache {
    struct Block { int device; void *p; size_t bytes; };
    std::mutex mu;
    std::vector<Block> blocks;   // oldest first
    size_t total = 0;
    size_t limit() {
        static const size_t v = [] {
            const char *e = std::getenv("RTB_BLOCK_CACHE_MB");
            return (size_t) (e ? std::max(0, std::atoi(e)) : 8192) << 20;
        }();
        return v;
    }
    // smallest kept block of this device that holds `bytes` without wasting more than half of itself
    cudaError_t acquire(int device, size_t bytes, void **out, size_t *got) {
        {
            std::lock_guard<std::mutex> lk(mu);
            int best = -1;
            for (int i = 0; i < (int) blocks.size(); ++i) {
                const Block &b = blocks[(size_t) i];
                if (b.device != device || b.bytes < bytes || b.bytes > bytes + bytes / 2 + (1u << 20)) continue;
                if (best < 0 || b.bytes < blocks[(size_t) best].bytes) best = i;
            }
            if (best >= 0) {
                *out = blocks[(size_t) best].p;
                *got = blocks[(size_t) best].bytes;
                total -= *got;
                blocks.erase(blocks.begin() + best);
                return cudaSuccess;
            }
        }
        *got = bytes;
        cudaError_t e = cudaMalloc(out, bytes);
        if (e != cudaSuccess && !defer_flag()) {   // out of memory with blocks parked here: give them back and try once more
            cudaGetLastError();
            trim(device);
            e = cudaMalloc(out, bytes);
        }
        // (a thread that defers its frees -- a rank thread of rtb_render_multi -- must not trim either: trim is cudaFree.  It
        // gets the error; rtb_render_multi hands the cache back once its threads have joined, so the caller's next frame finds
        // the memory)
        return e;
    }
    // Frees are DEFERRED on threads that asked for it (defer_frees): a thread that renders one device of a multi-device frame
    // must not call cudaFree -- it waits for a device to go idle, and a device whose NCCL kernel is waiting for THIS thread's
    // rank never does (seen as a hang of rtb_render_multi on 8 GPUs when a scratch block grew while the cache was full and the
    // oldest block belonged to another device).  The blocks go to `deferred` and are freed by flush_deferred() from a thread
    // that holds no rank (rtb_render_multi after its threads have joined, scene creation / destruction).
    std::vector<Block> deferred;
    static bool &defer_flag() {
        static thread_local bool f = false;
        return f;
    }
    void free_now(std::vector<Block> &drop, int restore_device) {
        for (const Block &b : drop) { cudaSetDevice(b.device); cudaFree(b.p); }
        if (!drop.empty() && restore_device >= 0) cudaSetDevice(restore_device);
        drop.clear();
    }
    void flush_deferred() {
        std::vector<Block> drop;
        {
            std::lock_guard<std::mutex> lk(mu);
            drop.swap(deferred);
        }
        if (drop.empty()) return;
        int cur = 0;
        cudaGetDevice(&cur);
        free_now(drop, cur);
    }
    // the caller guarantees that no queued work uses the block any more
    void release(int device, void *p, size_t bytes) {
        if (!p) return;
        std::vector<Block> drop;
        {
            std::lock_guard<std::mutex> lk(mu);
            if (bytes > limit()) drop.push_back(Block{device, p, bytes});
            else {
                blocks.push_back(Block{device, p, bytes});
                total += bytes;
                // at most 8 blocks per device; over the byte limit the oldest blocks go, this device's first
                auto count_dev = [&] { size_t c = 0; for (const Block &b : blocks) c += b.device == device; return c; };
                auto drop_oldest = [&](bool same_device_only) {
                    for (size_t i = 0; i < blocks.size(); ++i)
                        if (!same_device_only || blocks[i].device == device) {
                            drop.push_back(blocks[i]);
                            total -= blocks[i].bytes;
                            blocks.erase(blocks.begin() + (long) i);
                            return true;
                        }
                    return false;
                };
                while (count_dev() > 8) drop_oldest(true);
                while (total > limit() && (drop_oldest(true) || drop_oldest(false))) {}
            }
            if (defer_flag()) {
                deferred.insert(deferred.end(), drop.begin(), drop.end());
                drop.clear();
            }
        }
        free_now(drop, device);
    }
    void trim(int device) {   // device < 0: every device
        std::vector<Block> drop;
        {
            std::lock_guard<std::mutex> lk(mu);
            for (size_t i = 0; i < blocks.size();) {
                if (device < 0 || blocks[i].device == device) { drop.push_back(blocks[i]); total -= blocks[i].bytes; blocks.erase(blocks.begin() + (long) i); }
                else ++i;
            }
        }
        int cur = 0;
        cudaGetDevice(&cur);
        for (const Block &b : drop) { cudaSetDevice(b.device); cudaFree(b.p); }
        cudaSetDevice(cur);
    }
};
}  // namespace rtb
