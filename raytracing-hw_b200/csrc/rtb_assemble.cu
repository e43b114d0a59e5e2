// Kernels of the device-side scene assembly (rtb_assemble.cuh): one thread per post-build triangle writes its three
// records from the raw glTF views; the reference-order node array becomes the 64-byte two-child nodes level by level.
// HBM-bound byte work: coalesced record stores (4 + 4 + 3 float4 per thread), gathers through the index / vertex views.
#include "rtb_assemble.cuh"
#include "rtb_kernels.cuh"

namespace rtb {
namespace {

__global__ void __launch_bounds__(256) k_assemble_triangles(const DRawMesh *meshes, const uint32_t *mesh_first, int n_meshes, const uint32_t *tri_source,
                                                            uint32_t n_tris, float4 *tris, float4 *shade, float4 *tan, int *error) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_tris) return;
    const uint32_t src = tri_source[i];
    if (src >= mesh_first[n_meshes]) { atomicCAS(error, kAsmOk, kAsmSource); return; }
    const int m = mesh_of_source(mesh_first, n_meshes, src);
    float4 t[kTriStride], s[4], g[3];
    const int e = assemble_triangle_records(meshes[m], src - mesh_first[m], t, s, tan ? g : nullptr);
    if (e != kAsmOk) atomicCAS(error, kAsmOk, e);
#pragma unroll
    for (int k = 0; k < kTriStride; ++k) tris[(size_t) i * kTriStride + k] = t[k];
#pragma unroll
    for (int k = 0; k < 4; ++k) shade[(size_t) i * 4 + k] = s[k];
    if (tan) {
#pragma unroll
        for (int k = 0; k < 3; ++k) tan[(size_t) i * 3 + k] = g[k];
    }
}

// DTexture::quads of one texture: a thread per texel, coalesced 16-byte stores, the four taps gathered from two rows
__global__ void __launch_bounds__(256) k_texture_quads(const uchar4 *texels, uint4 *quads, int width, int height) {
    const size_t i = (size_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (size_t) width * (size_t) height) return;
    quads[i] = texture_quad(texels, width, height, (int) (i % (size_t) width), (int) (i / (size_t) width));
}

__global__ void __launch_bounds__(256) k_convert_level(const NodeItem *in, const unsigned *n_in_ptr, uint32_t n_nodes, uint32_t n_internal, uint32_t n_tris,
                                                       const float *aabb, const int32_t *links, BigLeafIndex big, DNode *nodes, NodeItem *out,
                                                       unsigned *n_out, int *error) {
    // (a frontier buffer holds n_internal items; after an overflow -- flagged in *error by the level before -- the counter
    // is larger than that)
    const unsigned n_in = min(*n_in_ptr, n_internal);
    for (unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < n_in; i += gridDim.x * blockDim.x) {
        NodeItem kids[2];
        int nk = 0;
        const int e = convert_node(in[i], n_nodes, n_internal, n_tris, aabb, links, big, nodes, kids, &nk);
        if (e != kAsmOk) { atomicCAS(error, kAsmOk, e); continue; }
        if (nk > 0) {
            const unsigned at = atomicAdd(n_out, (unsigned) nk);
            if (at + (unsigned) nk > n_internal) { atomicCAS(error, kAsmOk, kAsmNodeOrder); continue; }   // (more internal nodes than (n - 1) / 2: not a tree)
            for (int k = 0; k < nk; ++k) out[at + (unsigned) k] = kids[k];
        }
    }
}

__global__ void k_seed_root(NodeItem *items, unsigned *counts, uint32_t n_nodes) {
    items[0] = NodeItem{0u, 0u, 1u, n_nodes};
    counts[0] = 1u;
    counts[1] = 0u;
}

}  // namespace

// DFused::data of one material group: a thread per texel, 64 bytes stored per thread
__global__ void __launch_bounds__(256) k_fuse_textures(const uchar4 *base, const uchar4 *mr, const uchar4 *nrm, uint4 *out, int width, int height) {
    const size_t i = (size_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (size_t) width * (size_t) height) return;
    uint4 r[4];
    fused_texel(base, mr, nrm, width, height, (int) (i % (size_t) width), (int) (i / (size_t) width), r);
#pragma unroll
    for (int k = 0; k < 4; ++k) out[i * 4 + k] = r[k];
}
void launch_fuse_textures(const DFused &f, const uchar4 *base, const uchar4 *mr, const uchar4 *nrm, cudaStream_t st) {
    const size_t n = (size_t) f.width * (size_t) f.height;
    if (n == 0 || !f.data) return;
    k_fuse_textures<<<(unsigned) ((n + 255) / 256), 256, 0, st>>>(base, mr, nrm, const_cast<uint4 *>(f.data), f.width, f.height);
}

void launch_texture_quads(const DTexture &t, cudaStream_t st) {
    const size_t n = (size_t) t.width * (size_t) t.height;
    if (n == 0 || !t.quads) return;
    k_texture_quads<<<(unsigned) ((n + 255) / 256), 256, 0, st>>>(t.texels, const_cast<uint4 *>(t.quads), t.width, t.height);
}

void launch_assemble_triangles(const DRawMesh *d_meshes, const uint32_t *d_mesh_first, int n_meshes, const uint32_t *d_tri_source, uint32_t n_tris,
                               float4 *d_tris, float4 *d_shade, float4 *d_tan, int *d_error, cudaStream_t stream) {
    if (n_tris == 0) return;
    k_assemble_triangles<<<(n_tris + 255u) / 256u, 256, 0, stream>>>(d_meshes, d_mesh_first, n_meshes, d_tri_source, n_tris, d_tris, d_shade, d_tan, d_error);
}

// d_items: 2 x n_internal NodeItem (the two frontiers); d_counts: 2 unsigned; h_count: pinned or pageable host word.
// Returns the number of levels (= the depth of the internal-node tree), or -1 on a CUDA error.
int launch_convert_nodes(uint32_t n_nodes, uint32_t n_internal, uint32_t n_tris, const float *d_aabb, const int32_t *d_links, const BigLeafIndex &big,
                         DNode *d_nodes, NodeItem *d_items, unsigned *d_counts, int *d_error, cudaStream_t stream) {
    if (n_internal == 0) return 0;
    k_seed_root<<<1, 1, 0, stream>>>(d_items, d_counts, n_nodes);
    NodeItem *fr[2] = {d_items, d_items + n_internal};
    int levels = 0;
    unsigned n_in = 1;
    const int sm_grid = persistent_grid(256, 8);
    for (int cur = 0; n_in > 0; cur ^= 1) {
        ++levels;
        if (levels > 4096) return -1;
        const unsigned grid = (unsigned) std::min<size_t>(((size_t) n_in + 255) / 256, (size_t) sm_grid);
        k_convert_level<<<grid, 256, 0, stream>>>(fr[cur], d_counts + cur, n_nodes, n_internal, n_tris, d_aabb, d_links, big, d_nodes, fr[cur ^ 1],
                                                 d_counts + (cur ^ 1), d_error);
        // the next frontier's size decides the next launch: one 4-byte read-back per level (a tree has ~log2(n) + a few levels)
        unsigned next = 0;
        int flagged = 0;
        if (cudaMemcpyAsync(&next, d_counts + (cur ^ 1), sizeof(unsigned), cudaMemcpyDeviceToHost, stream) != cudaSuccess) return -1;
        if (cudaMemcpyAsync(&flagged, d_error, sizeof(int), cudaMemcpyDeviceToHost, stream) != cudaSuccess) return -1;
        if (cudaMemsetAsync(d_counts + cur, 0, sizeof(unsigned), stream) != cudaSuccess) return -1;
        if (cudaStreamSynchronize(stream) != cudaSuccess) return -1;
        if (flagged != 0) break;   // malformed input: the caller reads *d_error; no point in walking what is not a tree
        n_in = std::min(next, n_internal);
    }
    return levels;
}

}  // namespace rtb
