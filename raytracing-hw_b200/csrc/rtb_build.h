// GPU BVH builder with the reference's node AND triangle layout (BVH::buildBVH, src/core/bvh.cpp:5-175) -- the interface
// between the host builder (host/bvh_builder.cpp) and the kernels (csrc/rtb_build.cu).  Plain C++.
#pragma once
#include <cstddef>
#include <cstdint>
#include <string>
#include <vector>

namespace rtb {

enum : int32_t {
    kTopInternal = 0,   // left / right: indices into the node array returned with it
    kTopLeaf = 1,       // a leaf of the reference's tree: the split left one side empty (bvh.cpp:143-147); box valid
    kTopHost = 2,       // a range the GPU did not cut: [first, first + count) of the returned order is the range exactly as
                        // the reference's buildNode would find it -- the host builder carries on from there
};

struct GpuTopNode {
    float box[6];       // min xyz, max xyz (kTopInternal, kTopLeaf)
    int32_t left, right;
    int32_t split_dim;
    int32_t kind;
    uint32_t first, count;
};

struct GpuBuildStats {
    int levels = 0;
    long rounds = 0;              // partition rounds over all levels
    long heap_ranges = 0;         // introsort ranges that ran out of depth (heap-sorted, as libstdc++ does)
    long host_fallback_nodes = 0; // nodes of >= host_below triangles handed back uncut (no SAH improvement on a large node)
    double seconds = 0.0;         // everything: upload, kernels, download
    double kernel_seconds = 0.0;
};

// Builds every node of at least `host_below` triangles (>= 5) on GPU `device`.  tbox: n x 6 (Primitive::aabb()),
// tcenter: n x 3 (aabb().center()).  Returns false (err says why) when there is no device or a CUDA call failed; the
// caller then builds on the host.  depth_override >= 0 replaces introsort's 2 * lg(n) depth limit (tests).
bool gpu_build_top(const float *tbox, const float *tcenter, size_t n, size_t host_below, int device, int depth_override,
                   std::vector<GpuTopNode> &nodes, std::vector<uint32_t> &order, GpuBuildStats &stats, std::string &err);

// Test hook (tests/devsim/build_emul_check): the function the host builder calls for the GPU part -- null restores the real one.
// Lets the hand-over between the two builders (host/bvh_builder.cpp: adopt_all) run in a container without a GPU, against the
// builder's own source compiled for the host.
using GpuBuildFn = bool (*)(const float *, const float *, size_t, size_t, int, int, std::vector<GpuTopNode> &, std::vector<uint32_t> &,
                            GpuBuildStats &, std::string &);
void set_gpu_builder_for_tests(GpuBuildFn fn);

}  // namespace rtb
