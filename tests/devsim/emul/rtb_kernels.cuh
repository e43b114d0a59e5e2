// Test tooling: what csrc/rtb_dist.cu needs of csrc/rtb_kernels.cuh when it is compiled for the host (tests/devsim/dist_emul_check).
#pragma once
#include <cstddef>
#include <cstdint>
#include "cuda_emulation.h"
namespace rtb {
void launch_resolve(const float *accum, size_t n_pixels, int samples, uint8_t *rgb8, cudaStream_t st);
}
