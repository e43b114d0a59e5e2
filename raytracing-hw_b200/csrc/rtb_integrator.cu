// Path-tracing megakernel (sm_100a): replaces the two loops of Scene::render (src/core/scene.cpp:31-64).
//
// One thread owns one pixel and runs all of its samples back to back, because the reference draws every sample of a
// pixel from ONE sequential minstd_rand stream whose consumption is data dependent (rejection loops, early exits):
// sample s cannot start before sample s-1 has finished if the stream is to be reproduced.  A warp covers an 8x4
// pixel tile (primary rays and first bounces of neighbouring pixels touch the same nodes), a block two such tiles.
// Traversal stack and the per-bounce (emission, weight) stack live in local memory (L1-resident).
// Compiled with --fmad=false: see rtb_device.cuh / rtb_shade.cuh for the arithmetic contract.
#include "rtb_kernels.cuh"
#include "rtb_shade.cuh"

namespace rtb {

constexpr int kTileW = 8, kTileH = 4, kWarpsPerBlock = 2;

template <int CULL, bool COUNT>
__global__ void __launch_bounds__(32 * kWarpsPerBlock) k_render(const __grid_constant__ DScene sc, const RenderLaunch p) {
    const int tiles_x = (p.width + kTileW - 1) / kTileW;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const long long tile = (long long) blockIdx.x * kWarpsPerBlock + warp;
    const int tx = (int) (tile % tiles_x), ty = (int) (tile / tiles_x);
    const int px = tx * kTileW + (lane & (kTileW - 1));
    const int lr = ty * kTileH + (lane >> 3);   // local (output) row
    int py = p.row_begin + lr;
    if (p.il_count > 1) py = p.row_begin + ((lr / p.il_rows) * p.il_count + p.il_index) * p.il_rows + lr % p.il_rows;
    PathCounters pc = {0ull, 0ull, {0ull, 0ull}, {0ull, 0ull}};
    if (px < p.width && lr < p.local_rows && py < p.row_end) {
        const V3 sum = render_pixel<CULL, COUNT>(sc, px, py, p.width, p.height, p.samples, p.pixel0_seed, &pc);
        const size_t o = ((size_t) lr * p.width + px) * 3;
        if (p.d_accum) {
            p.d_accum[o] = sum.x;
            p.d_accum[o + 1] = sum.y;
            p.d_accum[o + 2] = sum.z;
        }
        if (p.d_rgb8) {
            const float normalizer = 1.f / (float) p.samples;
            p.d_rgb8[o] = resolve_channel(sum.x, normalizer);
            p.d_rgb8[o + 1] = resolve_channel(sum.y, normalizer);
            p.d_rgb8[o + 2] = resolve_channel(sum.z, normalizer);
        }
    }
    if (p.d_counters) {
        unsigned long long c[6] = {pc.rays, pc.light_queries, pc.tests.box + pc.light.box, pc.tests.tri + pc.light.tri, pc.light.box, pc.light.tri};
#pragma unroll
        for (int k = 0; k < 6; ++k) {
            for (int o = 16; o > 0; o >>= 1) c[k] += __shfl_down_sync(0xffffffffu, c[k], o);
            if (lane == 0 && c[k]) atomicAdd(&p.d_counters[k], c[k]);
        }
    }
}

void launch_render(const DScene &sc, const RenderLaunch &p, cudaStream_t stream) {
    const int rows = p.local_rows;
    if (rows <= 0 || p.width <= 0 || p.samples <= 0) return;
    const long long tiles = (long long) ((p.width + kTileW - 1) / kTileW) * ((rows + kTileH - 1) / kTileH);
    const unsigned grid = (unsigned) ((tiles + kWarpsPerBlock - 1) / kWarpsPerBlock);
    const unsigned block = 32 * kWarpsPerBlock;
    if (p.cull_mode == 0) {
        if (p.count_tests) k_render<0, true><<<grid, block, 0, stream>>>(sc, p);
        else k_render<0, false><<<grid, block, 0, stream>>>(sc, p);
    } else {
        if (p.count_tests) k_render<1, true><<<grid, block, 0, stream>>>(sc, p);
        else k_render<1, false><<<grid, block, 0, stream>>>(sc, p);
    }
}

// The resolve loop alone (src/core/scene.cpp:54-63), for accumulation buffers combined across GPUs.
__global__ void __launch_bounds__(256) k_resolve(const float *__restrict__ accum, size_t n_values, float normalizer, uint8_t *__restrict__ rgb8) {
    const size_t i = (size_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n_values) rgb8[i] = resolve_channel(accum[i], normalizer);
}

void launch_resolve(const float *d_accum, size_t n_pixels, int samples, uint8_t *d_rgb8, cudaStream_t stream) {
    const size_t n = n_pixels * 3;
    if (n == 0) return;
    k_resolve<<<(unsigned) ((n + 255) / 256), 256, 0, stream>>>(d_accum, n, 1.f / (float) samples, d_rgb8);
}

}  // namespace rtb
