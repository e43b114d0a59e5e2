// Batched closest-hit and light-pdf kernels (sm_100a).  One ray per thread, per-thread traversal stack in local
// memory, node = 4 x LDG.128, triangle = 3 x LDG.128 through the read-only path.
// Compiled with --fmad=false (decision-exact float arithmetic, see rtb_device.cuh).
#include "rtb_kernels.cuh"
#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <mutex>
#include "rtb_shade.cuh"
#include "rtb_trace.cuh"

namespace rtb {

// Replaces BVH::intersect (src/core/bvh.cpp:239-243) for a batch.
template <int CULL, bool COUNT>
__global__ void __launch_bounds__(128) k_closest_hit(const __grid_constant__ DScene sc, const float *__restrict__ rays6, size_t n,
                                                     int *__restrict__ id, float *__restrict__ t, float *__restrict__ u,
                                                     float *__restrict__ v, uint8_t *__restrict__ inside,
                                                     unsigned long long *__restrict__ counters) {
    const size_t i = (size_t) blockIdx.x * blockDim.x + threadIdx.x;
    Counters cnt = {0ull, 0ull};
    if (i < n) {
        const float2 *rp = reinterpret_cast<const float2 *>(rays6 + i * 6);
        const float2 r0 = __ldg(rp), r1 = __ldg(rp + 1), r2 = __ldg(rp + 2);
        Ray r;
        make_ray(r, r0.x, r0.y, r1.x, r1.y, r2.x, r2.y);
        Hit h;
        const bool ok = closest_hit<CULL, COUNT>(sc.bvh, r, h, &cnt);
        id[i] = ok ? h.id : -1;
        if (t) t[i] = ok ? h.t : 0.f;
        if (u) u[i] = ok ? h.u : 0.f;
        if (v) v[i] = ok ? h.v : 0.f;
        if (inside) {
            uint8_t in = 0;
            if (ok) {
                // `if (dot(ray.direction, normal) > 0) inside = true` (primitive.cpp:51-54)
                const float4 q2 = __ldg(sc.bvh.tris + (size_t) h.id * kTriStride + 2);
                float dn = r.d[0] * q2.y;
                dn = dn + r.d[1] * q2.z;
                dn = dn + r.d[2] * q2.w;
                in = dn > 0.f ? 1 : 0;
            }
            inside[i] = in;
        }
    }
    if (COUNT) {
        // warp-aggregate, then one atomic per warp and counter
        for (int o = 16; o > 0; o >>= 1) {
            cnt.box += __shfl_down_sync(0xffffffffu, cnt.box, o);
            cnt.tri += __shfl_down_sync(0xffffffffu, cnt.tri, o);
        }
        if ((threadIdx.x & 31) == 0) {
            atomicAdd(&counters[0], cnt.box);
            atomicAdd(&counters[1], cnt.tri);
        }
    }
}

// Persistent ray-queue variant (RTB_CULL_GLOBAL): see rtb_trace.cuh.  grid = a few CTAs per SM; lanes pull rays.
template <bool COUNT, int MINB, int MODE>
__global__ void __launch_bounds__(128, MINB) k_closest_hit_queue(const __grid_constant__ DScene sc, const float *__restrict__ rays6, unsigned n,
                                                           unsigned *__restrict__ queue_head, int *__restrict__ id,
                                                           float *__restrict__ t, float *__restrict__ u, float *__restrict__ v,
                                                           uint8_t *__restrict__ inside, unsigned long long *__restrict__ counters,
                                                           const TracePolicy pol) {
    QueueCounters cnt = {0ull, 0ull, 0u};
    trace_persistent<COUNT, MODE, RTB_RING_BATCH>(
        sc.bvh, n, queue_head,
        [&](unsigned i, LaneTraversal<COUNT, RTB_RING_BATCH> &L, unsigned &tag) {
            const float2 *rp = reinterpret_cast<const float2 *>(rays6 + (size_t) i * 6);
            const float2 r0 = __ldg(rp), r1 = __ldg(rp + 1), r2 = __ldg(rp + 2);
            make_ray(L.r, r0.x, r0.y, r1.x, r1.y, r2.x, r2.y);
            tag = i;
            return 1;
        },
        [&](unsigned i, const Ray &r, const Hit &h) {
            const bool ok = h.id >= 0;
            id[i] = ok ? h.id : -1;
            if (t) t[i] = ok ? h.t : 0.f;
            if (u) u[i] = ok ? h.u : 0.f;
            if (v) v[i] = ok ? h.v : 0.f;
            if (inside) {
                uint8_t in = 0;
                if (ok) {   // `if (dot(ray.direction, normal) > 0) inside = true` (primitive.cpp:51-54)
                    const float4 q2 = __ldg(sc.bvh.tris + (size_t) h.id * kTriStride + 2);
                    float dn = r.d[0] * q2.y;
                    dn = dn + r.d[1] * q2.z;
                    dn = dn + r.d[2] * q2.w;
                    in = dn > 0.f ? 1 : 0;
                }
                inside[i] = in;
            }
        },
        [](unsigned, LaneTraversal<COUNT, RTB_RING_BATCH> &) {}, 0,   // a batch is one launch: nothing is ever suspended
        &cnt, pol);
    if (COUNT) {
        for (int o = 16; o > 0; o >>= 1) {
            cnt.box += __shfl_down_sync(0xffffffffu, cnt.box, o);
            cnt.tri += __shfl_down_sync(0xffffffffu, cnt.tri, o);
        }
        if ((threadIdx.x & 31) == 0) {
            atomicAdd(&counters[0], cnt.box);
            atomicAdd(&counters[1], cnt.tri);
        }
    }
}

TracePolicy trace_policy() {
    static TracePolicy pol = [] {
        TracePolicy p = {4, 1, 1};   // profiles/r01f_closest_hit_mode_blocks_sweep.jsonl
        if (const char *e = std::getenv("RTB_TRACE_POLICY")) {
            int a = 0, b = 0, c = 0;
            if (std::sscanf(e, "%d,%d,%d", &a, &b, &c) == 3 && a >= 1 && a <= 32 && b >= 1 && c >= 1) p = {a, b, c};
        }
        return p;
    }();
    return pol;
}

int trace_mode() {
    static int v = [] {
        int m = kDefaultTraceMode;
        if (const char *e = std::getenv("RTB_TRACE_MODE")) m = std::max(3, std::min(5, std::atoi(e)));
        return m;
    }();
    return v;
}

int persistent_blocks_per_sm() {
    static int v = [] {
        int b = kDefaultTraceBlocks;
        if (const char *e = std::getenv("RTB_TRACE_BLOCKS")) b = std::max(1, std::min(10, std::atoi(e)));
        return b;
    }();
    return v;
}

int persistent_grid(int threads_per_block, int blocks_per_sm) {
    static int sms = 0;
    if (sms == 0) {
        int dev = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
        if (sms <= 0) sms = 148;
    }
    (void) threads_per_block;
    return sms * blocks_per_sm;
}

void launch_closest_hit(const DScene &sc, const float *d_rays6, size_t n, int cull_mode, int *d_id, float *d_t, float *d_u,
                        float *d_v, uint8_t *d_inside, unsigned long long *d_counters, unsigned *d_queue_head, cudaStream_t stream) {
    if (n == 0) return;
    const unsigned block = 128;
    if (cull_mode == 1 && d_queue_head != nullptr && n < 0xFFFFFF00u) {
        cudaMemsetAsync(d_queue_head, 0, sizeof(unsigned), stream);
        const int bps = persistent_blocks_per_sm();
        const unsigned pg = (unsigned) std::min<size_t>((n + block - 1) / block, (size_t) persistent_grid(block, bps));
#define RTB_LAUNCH_Q(COUNT, MINB, MODE) k_closest_hit_queue<COUNT, MINB, MODE><<<pg, block, 0, stream>>>(sc, d_rays6, (unsigned) n, d_queue_head, d_id, d_t, d_u, d_v, d_inside, d_counters, trace_policy())
#define RTB_LAUNCH_QM(MODE)                                  \
        if (d_counters) RTB_LAUNCH_Q(true, 10, MODE);        \
        else if (bps <= 9) RTB_LAUNCH_Q(false, 9, MODE);     \
        else RTB_LAUNCH_Q(false, 10, MODE)
        if (trace_mode() == 3) { RTB_LAUNCH_QM(3); } else if (trace_mode() == 5) { RTB_LAUNCH_QM(5); } else { RTB_LAUNCH_QM(4); }
#undef RTB_LAUNCH_QM
#undef RTB_LAUNCH_Q
        return;
    }
    const unsigned grid = (unsigned) ((n + block - 1) / block);
    const bool count = d_counters != nullptr;
    if (cull_mode == 0) {
        if (count) k_closest_hit<0, true><<<grid, block, 0, stream>>>(sc, d_rays6, n, d_id, d_t, d_u, d_v, d_inside, d_counters);
        else k_closest_hit<0, false><<<grid, block, 0, stream>>>(sc, d_rays6, n, d_id, d_t, d_u, d_v, d_inside, d_counters);
    } else {
        if (count) k_closest_hit<1, true><<<grid, block, 0, stream>>>(sc, d_rays6, n, d_id, d_t, d_u, d_v, d_inside, d_counters);
        else k_closest_hit<1, false><<<grid, block, 0, stream>>>(sc, d_rays6, n, d_id, d_t, d_u, d_v, d_inside, d_counters);
    }
}

// ---- closest hit for callers that must not allocate (rtb_closest_hit_static, include/rtb.h) --------------------------
// The persistent queue kernel needs one word of global memory besides its inputs and outputs: the cursor its warps
// advance to claim rays.  A scene owns cudaMalloc'ed cursors; this variant takes the cursor from a table that is part
// of the library image instead, so the call path contains no allocation at all.  64 cursors, one per 256-byte line
// (16 KB on every GPU that loads librtb.so, until the process exits), handed out round-robin per device; a cursor is
// cleared by a stream-ordered memset in front of the launch that uses it, because the table keeps whatever the last
// call left in it.
constexpr unsigned kStaticCursors = 64;
constexpr unsigned kStaticCursorStride = 256 / sizeof(unsigned);
constexpr int kStaticMaxDevices = 64;
static __device__ __align__(256) unsigned g_static_cursors[kStaticCursors * kStaticCursorStride];

// `device` must be the current device.  In host code `g_static_cursors` is only a handle for the runtime, not an address
// a kernel could use: the table's address on THIS device is asked for once (cudaGetSymbolAddress) and remembered.
cudaError_t launch_closest_hit_static(const DScene &sc, int device, const float *d_rays6, size_t n, int cull_mode, int *d_id, float *d_t,
                                      float *d_u, float *d_v, uint8_t *d_inside, unsigned long long *d_counters, cudaStream_t stream) {
    static std::mutex mu;
    static unsigned *table[kStaticMaxDevices] = {};
    static unsigned next[kStaticMaxDevices] = {};
    if (device < 0 || device >= kStaticMaxDevices) return cudaErrorInvalidDevice;
    unsigned *cursor = nullptr;
    {
        std::lock_guard<std::mutex> lock(mu);
        if (!table[device]) {
            void *p = nullptr;
            const cudaError_t e = cudaGetSymbolAddress(&p, g_static_cursors);
            if (e != cudaSuccess) return e;
            table[device] = static_cast<unsigned *>(p);
        }
        cursor = table[device] + (size_t) (next[device]++ % kStaticCursors) * kStaticCursorStride;
    }
    // (launch_closest_hit clears the cursor on `stream` before the queue kernel; RTB_CULL_REFERENCE runs the
    // one-ray-per-thread kernel, which has no cursor)
    launch_closest_hit(sc, d_rays6, n, cull_mode, d_id, d_t, d_u, d_v, d_inside, d_counters, cursor, stream);
    return cudaGetLastError();
}

// ManyLightsDistribution::pdf (src/utils/random.cpp:179-188) for a batch of (point, direction).
__global__ void __launch_bounds__(128) k_light_pdf(const __grid_constant__ DScene sc, const float *__restrict__ rays6, size_t n,
                                                   float *__restrict__ pdf) {
    const size_t i = (size_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float *rp = rays6 + i * 6;
    float prob = 0.f;
    if (sc.n_lights > 0) prob = pdf_lights<false>(sc, V3{rp[0], rp[1], rp[2]}, V3{rp[3], rp[4], rp[5]}, nullptr);
    pdf[i] = prob;
}

// ---- L2 bandwidth probes: the denominators the L2-resident scenes' traversal is compared with (SURVEY.md section 8d asks
// for a measured L2 figure).  `stream`: every thread walks the buffer with coalesced 16-byte loads, `reps` passes over
// a buffer that fits the L2.  `gather64`: every thread reads 64-byte records at pseudo-random indices with two 256-bit
// loads -- the access pattern of an incoherent ray's node / triangle fetch, without any arithmetic behind it.
__global__ void __launch_bounds__(256) k_l2_stream(const float4 *__restrict__ p, size_t n4, int reps, float *__restrict__ out) {
    float acc = 0.f;
    const size_t stride = (size_t) gridDim.x * blockDim.x;
    for (int r = 0; r < reps; ++r)
        for (size_t i = (size_t) blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += stride) {
            const float4 v = __ldcg(p + i);   // (past the L1: one SM's share of the buffer would fit it)
            acc += v.x + v.y + v.z + v.w;
        }
    if (acc == 123.456f) out[0] = acc;   // (keeps the loads alive)
}
__global__ void __launch_bounds__(256) k_l2_gather64(const float4 *__restrict__ p, unsigned n_rec, int iters, float *__restrict__ out) {
    float acc = 0.f;
    unsigned x = (blockIdx.x * blockDim.x + threadIdx.x) * 2654435761u + 12345u;
    for (int it = 0; it < iters; ++it) {
        x = x * 1664525u + 1013904223u;
        const float4 *q = p + (size_t) ((x >> 8) % n_rec) * 4;
        float4 a, b, c;
        int4 d;
        ldg256(q, a, b);
        ldg256(q + 2, c, d);
        acc += a.x + b.y + c.z + (float) d.w;
    }
    if (acc == 123.456f) out[0] = acc;
}
int measure_l2(size_t bytes, double *stream_gbs, double *gather64_gbs) {
    bytes = std::max<size_t>(bytes / 64 * 64, 1 << 20);
    float4 *buf = nullptr;
    float *out = nullptr;
    if (cudaMalloc(&buf, bytes) != cudaSuccess || cudaMalloc(&out, 4) != cudaSuccess) { cudaGetLastError(); if (buf) cudaFree(buf); return 1; }
    cudaMemset(buf, 0, bytes);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    const int grid = persistent_grid(256, 8);
    const int reps = 200, iters = 4000;
    float ms = 0.f;
    k_l2_stream<<<grid, 256>>>(buf, bytes / 16, 2, out);   // warm the L2
    cudaEventRecord(e0);
    k_l2_stream<<<grid, 256>>>(buf, bytes / 16, reps, out);
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    cudaEventElapsedTime(&ms, e0, e1);
    if (stream_gbs) *stream_gbs = (double) bytes * reps / (ms * 1e-3) * 1e-9;
    cudaEventRecord(e0);
    k_l2_gather64<<<grid, 256>>>(buf, (unsigned) (bytes / 64), iters, out);
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    cudaEventElapsedTime(&ms, e0, e1);
    if (gather64_gbs) *gather64_gbs = (double) grid * 256 * iters * 64.0 / (ms * 1e-3) * 1e-9;
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    cudaFree(buf);
    cudaFree(out);
    return cudaGetLastError() == cudaSuccess ? 0 : 1;
}

void launch_light_pdf(const DScene &sc, const float *d_rays6, size_t n, float *d_pdf, cudaStream_t stream) {
    if (n == 0) return;
    const unsigned block = 128;
    k_light_pdf<<<(unsigned) ((n + block - 1) / block), block, 0, stream>>>(sc, d_rays6, n, d_pdf);
}

}  // namespace rtb
