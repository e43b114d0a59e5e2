// Test tooling: the few CUDA names csrc/rtb_build.cu uses, for a HOST compile of that file (tests/devsim/build_emul): device
// memory is host memory, a kernel launch runs the kernel's threads one after the other, atomics are plain operations.  The
// block-wide scans are the one thing a sequential run of threads cannot do; rtb_build.cu swaps them for an in-order loop with the
// same operator under RTB_BUILD_EMULATE.  Nothing here is linked into librtb.so.
#pragma once
#include <cstdint>
#include <cstdlib>
#include <cstring>

#define __global__
#define __device__
#define __host__
#define __forceinline__ inline
#define __launch_bounds__(...)
#define __restrict__

struct EmuDim { unsigned x = 0, y = 0, z = 0; };
inline EmuDim threadIdx, blockIdx, blockDim, gridDim;

typedef int cudaError_t;
typedef void *cudaStream_t;
enum { cudaSuccess = 0 };
enum { cudaMemcpyHostToDevice = 1, cudaMemcpyDeviceToHost = 2, cudaMemcpyDeviceToDevice = 3 };
enum { cudaStreamNonBlocking = 1 };
inline cudaError_t cudaGetDeviceCount(int *n) { *n = 1; return cudaSuccess; }
inline thread_local int emu_device = 0;    // the current device is per thread, as in CUDA
inline int emu_frees[64] = {};             // cudaFree calls by current device
inline void (*emu_free_hook)() = nullptr;  // called by every cudaFree of a non-null pointer (tests: who frees, and when)
inline cudaError_t cudaGetDevice(int *d) { *d = emu_device; return cudaSuccess; }
inline int emu_bad_device = -1;            // tests: cudaSetDevice(this device) fails
inline cudaError_t cudaSetDevice(int d) { if (d == emu_bad_device) return 101; emu_device = d; return cudaSuccess; }
inline cudaError_t cudaGetLastError() { return cudaSuccess; }
inline const char *cudaGetErrorString(cudaError_t) { return "emulated"; }
inline size_t emu_malloc_limit = ~(size_t) 0;   // tests: requests above this many bytes fail while emu_live_bytes-free accounting is off
inline int emu_malloc_fail_next = 0;            // tests: the next k cudaMalloc calls report out of memory
inline cudaError_t cudaMalloc(void **p, size_t bytes) {
    if (bytes > emu_malloc_limit || emu_malloc_fail_next > 0) {
        if (emu_malloc_fail_next > 0) --emu_malloc_fail_next;
        *p = nullptr;
        return 2;
    }
    *p = std::malloc(bytes);
    return *p ? cudaSuccess : 2;
}
inline cudaError_t cudaFree(void *p) {
    if (p) {
        __sync_fetch_and_add(&emu_frees[emu_device & 63], 1);
        if (emu_free_hook) emu_free_hook();
    }
    std::free(p);
    return cudaSuccess;
}
inline cudaError_t cudaMemsetAsync(void *p, int v, size_t bytes, cudaStream_t) { std::memset(p, v, bytes); return cudaSuccess; }
typedef void *cudaEvent_t;
inline cudaError_t cudaEventCreate(cudaEvent_t *e) { *e = nullptr; return cudaSuccess; }
inline cudaError_t cudaEventRecord(cudaEvent_t, cudaStream_t) { return cudaSuccess; }
inline cudaError_t cudaEventDestroy(cudaEvent_t) { return cudaSuccess; }
inline cudaError_t cudaEventElapsedTime(float *ms, cudaEvent_t, cudaEvent_t) { *ms = 0.f; return cudaSuccess; }
inline cudaError_t cudaMemcpyAsync(void *dst, const void *src, size_t bytes, int, cudaStream_t) { std::memcpy(dst, src, bytes); return cudaSuccess; }
inline cudaError_t cudaStreamSynchronize(cudaStream_t) { return cudaSuccess; }
inline cudaError_t cudaStreamCreateWithFlags(cudaStream_t *s, int) { *s = nullptr; return cudaSuccess; }
inline cudaError_t cudaStreamDestroy(cudaStream_t) { return cudaSuccess; }

inline uint32_t atomicAdd(uint32_t *p, uint32_t v) { const uint32_t o = *p; *p = o + v; return o; }
inline unsigned long long atomicMin(unsigned long long *p, unsigned long long v) { const unsigned long long o = *p; if (v < o) *p = v; return o; }
// one thread at a time: a "warp" never agrees, so the per-lane path of a warp-aggregated atomic is the one that runs
template <class T> inline T __shfl_sync(unsigned, T v, int) { return v; }
template <class T> inline T __shfl_down_sync(unsigned, T v, int) { return v; }
inline bool __all_sync(unsigned, bool) { return false; }

struct EmuCfg { unsigned grid; unsigned block; };
inline EmuCfg emu_cfg(unsigned grid, int block, int, cudaStream_t) { return EmuCfg{grid, (unsigned) block}; }
// kernel<<<grid, block, 0, stream>>>(args)  ->  EMU_LAUNCH(kernel, (args), grid, block, 0, stream)   (tests/devsim/Makefile: sed)
#define EMU_LAUNCH(kernel, args, ...)                                        \
    do {                                                                     \
        const EmuCfg c_ = emu_cfg(__VA_ARGS__);                              \
        blockDim.x = c_.block;                                               \
        gridDim.x = c_.grid;                                                 \
        for (unsigned b_ = 0; b_ < c_.grid; ++b_)                            \
            for (unsigned t_ = 0; t_ < c_.block; ++t_) {                     \
                blockIdx.x = b_;                                             \
                threadIdx.x = t_;                                            \
                kernel args;                                                 \
            }                                                                \
    } while (0)
