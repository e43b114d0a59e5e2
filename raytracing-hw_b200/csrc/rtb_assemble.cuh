// Device-side scene assembly: what the reference does on the host while it parses the
// file -- index fetch, world transform, p0 / p1-p0 / p2-p0, normal transform + normalise, texcoords, tangents
// (src/io/scene_parser.cpp:280-345), the lazily cached geometric normal (src/geometry/primitive.cpp:77-84) -- and the
// conversion of the reference-order Node array (src/core/bvh.h:10-17) into the 64-byte two-child nodes, as per-triangle /
// per-node functions.  rtb_assemble.cu runs them in kernels over the raw glTF views; tests/devsim runs the same source on
// the host.  Every buffer they produce is byte-equal to what rtb_layout.hpp builds from the host-assembled arrays
// (tests/test_assembly.py, tests/test_gpu_assembly.py): each expression keeps the host's operation order and rounds
// after every operation (explicit _rn intrinsics on the device, so no contraction whatever the build flags are).
#pragma once
#include "../../include/rtb.h"
#include "rtb_device.cuh"

namespace rtb {

// one glTF primitive as the file delivers it; the views point at the uploaded bytes (device) / the caller's arrays (host)
struct DRawMesh {
    const uint8_t *idx, *pos, *nrm, *tan, *uv;
    uint32_t first_tri, n_tris, n_vertices;
    int index_bytes, idx_stride, pos_stride, nrm_stride, tan_stride, uv_stride;
    int normal_mode, flags, mesh_id, pad_;
    double transform[16], normal_transform[16];
};

// errors a kernel can flag (first one wins)
enum AssembleError : int {
    kAsmOk = 0,
    kAsmVertexIndex = 1,      // "[gltf check] vertex index out of range"
    kAsmSource = 2,           // tri_source entry beyond the meshes' triangles
    kAsmNodeOrder = 3,        // nodes are not in the reference's order (desc(X) = [L, R] + desc(R) + desc(L))
    kAsmLeafRange = 4,        // "leaf range out of bounds"
    kAsmChildren = 5,         // "internal node without two children"
    kAsmSplitDim = 6,         // "internal node with invalid split_dim"
    kAsmBigLeaf = 7,          // a leaf of more than 7 triangles missing from the host's table
};

namespace asmath {
RTB_HD float fmul(float a, float b) {
#ifdef __CUDA_ARCH__
    return __fmul_rn(a, b);
#else
    return a * b;
#endif
}
RTB_HD float fadd(float a, float b) {
#ifdef __CUDA_ARCH__
    return __fadd_rn(a, b);
#else
    return a + b;
#endif
}
RTB_HD float fsub(float a, float b) {
#ifdef __CUDA_ARCH__
    return __fsub_rn(a, b);
#else
    return a - b;
#endif
}
RTB_HD float fsqrt(float a) {
#ifdef __CUDA_ARCH__
    return __fsqrt_rn(a);
#else
    return std::sqrt(a);
#endif
}
RTB_HD float frcp(float a) {   // 1.f / a
#ifdef __CUDA_ARCH__
    return __fdiv_rn(1.f, a);
#else
    return 1.f / a;
#endif
}
// multiply(matrix4<double>, vector4f) (src/utils/matrix.h:66-73): double products, FLOAT accumulator; component i
RTB_HD float mat_vec4_component(const double *m, const float t[4], int i) {
    float acc = 0.f;
    for (int j = 0; j < 4; ++j) {
#ifdef __CUDA_ARCH__
        acc = __double2float_rn(__dadd_rn((double) acc, __dmul_rn(m[4 * j + i], (double) t[j])));
#else
        acc = (float) ((double) acc + m[4 * j + i] * (double) t[j]);
#endif
    }
    return acc;
}
// normal(vector3f) (src/utils/vector.h:169-174)
RTB_HD void normalize3(float v[3]) {
    float mod = 0.f;
    for (int i = 0; i < 3; ++i) mod = fadd(mod, fmul(v[i], v[i]));
    mod = fsqrt(mod);
    mod = frcp(mod);
    for (int i = 0; i < 3; ++i) v[i] = fmul(v[i], mod);
}
RTB_HD void cross3(const float *U, const float *V, float n[3]) {   // vector.h:372-378
    n[0] = fsub(fmul(U[1], V[2]), fmul(V[1], U[2]));
    n[1] = fsub(fmul(U[2], V[0]), fmul(V[2], U[0]));
    n[2] = fsub(fmul(U[0], V[1]), fmul(V[0], U[1]));
}
RTB_HD float load_f(const uint8_t *p) { return *reinterpret_cast<const float *>(p); }
}  // namespace asmath

// the index fetch of scene_parser.cpp:285-304
RTB_HD uint32_t raw_vertex(const DRawMesh &m, uint32_t tri, int v) {
    const size_t k = (size_t) tri * 3 + (size_t) v;
    if (!m.idx) return (uint32_t) k;
    const uint8_t *p = m.idx + k * (size_t) m.idx_stride;
    if (m.index_bytes == 2) return *reinterpret_cast<const uint16_t *>(p);
    return *reinterpret_cast<const uint32_t *>(p);
}

// One triangle of a raw mesh -> its traversal record (kTriStride float4: p0, U, V, Ng), its shading record (4 float4:
// normals, texcoords, mesh id) and, when tan_rec != null, its tangent record (3 float4).  Returns kAsmOk or an error.
RTB_HD int assemble_triangle_records(const DRawMesh &m, uint32_t tri, float4 *tri_rec, float4 *shade_rec, float4 *tan_rec) {
    using namespace asmath;
    uint32_t index[3];
    int err = kAsmOk;
    for (int v = 0; v < 3; ++v) {
        index[v] = raw_vertex(m, tri, v);
        if (index[v] >= m.n_vertices) { index[v] = 0; err = kAsmVertexIndex; }
    }
    float P[3][3];
    for (int v = 0; v < 3; ++v) {
        const uint8_t *p = m.pos + (size_t) index[v] * (size_t) m.pos_stride;
        const float in[4] = {load_f(p), load_f(p + 4), load_f(p + 8), 1.f};
        for (int k = 0; k < 3; ++k) P[v][k] = (m.flags & RTB_RAW_WORLD_POSITIONS) ? in[k] : mat_vec4_component(m.transform, in, k);
    }
    float U[3], V[3], Ng[3];
    for (int k = 0; k < 3; ++k) {
        U[k] = fadd(P[1][k], -P[0][k]);   // operator-(a,b) = a + (-b) (vector.h:326-328)
        V[k] = fadd(P[2][k], -P[0][k]);
    }
    cross3(U, V, Ng);                     // get_geometric_normal (primitive.cpp:77-84)
    float flat[3] = {Ng[0], Ng[1], Ng[2]};
    normalize3(Ng);
    tri_rec[0] = make_float4(P[0][0], P[0][1], P[0][2], U[0]);
    tri_rec[1] = make_float4(U[1], U[2], V[0], V[1]);
    tri_rec[2] = make_float4(V[2], Ng[0], Ng[1], Ng[2]);
    tri_rec[3] = make_float4(0.f, 0.f, 0.f, 0.f);

    float nn[9], uv[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    if (m.normal_mode == RTB_NORMALS_FLAT) {
        normalize3(flat);
        for (int v = 0; v < 3; ++v)
            for (int k = 0; k < 3; ++k) nn[v * 3 + k] = flat[k];
    } else if (m.normal_mode == RTB_NORMALS_VERBATIM) {
        for (int v = 0; v < 3; ++v) {
            const uint8_t *p = m.nrm + (size_t) index[v] * (size_t) m.nrm_stride;
            for (int k = 0; k < 3; ++k) nn[v * 3 + k] = load_f(p + 4 * k);
        }
    } else {   // scene_parser.cpp:316-325
        for (int v = 0; v < 3; ++v) {
            float in[4] = {0.f, 0.f, 1.f, 0.f};
            if (m.nrm) {
                const uint8_t *p = m.nrm + (size_t) index[v] * (size_t) m.nrm_stride;
                in[0] = load_f(p); in[1] = load_f(p + 4); in[2] = load_f(p + 8);
            }
            float res[3];
            for (int k = 0; k < 3; ++k) res[k] = mat_vec4_component(m.normal_transform, in, k);
            normalize3(res);
            for (int k = 0; k < 3; ++k) nn[v * 3 + k] = res[k];
        }
    }
    if (m.uv)
        for (int v = 0; v < 3; ++v) {
            const uint8_t *p = m.uv + (size_t) index[v] * (size_t) m.uv_stride;
            uv[v * 2] = load_f(p);
            uv[v * 2 + 1] = load_f(p + 4);
        }
    shade_rec[0] = make_float4(nn[0], nn[1], nn[2], nn[3]);
    shade_rec[1] = make_float4(nn[4], nn[5], nn[6], nn[7]);
    shade_rec[2] = make_float4(nn[8], uv[0], uv[1], uv[2]);
    shade_rec[3] = make_float4(uv[3], uv[4], uv[5], u2f((uint32_t) m.mesh_id));
    if (tan_rec) {
        for (int v = 0; v < 3; ++v) {
            float4 t = make_float4(0.f, 0.f, 0.f, 0.f);
            if (m.tan) {
                const uint8_t *p = m.tan + (size_t) index[v] * (size_t) m.tan_stride;
                t = make_float4(load_f(p), load_f(p + 4), load_f(p + 8), load_f(p + 12));
            }
            tan_rec[v] = t;
        }
    }
    return err;
}

// mesh of loader-order triangle `src`: first[] = n_meshes + 1 running triangle counts
RTB_HD int mesh_of_source(const uint32_t *first, int n_meshes, uint32_t src) {
    int lo = 0, hi = n_meshes;   // first[lo] <= src < first[hi]
    while (hi - lo > 1) {
        const int mid = (lo + hi) >> 1;
        if (first[mid] <= src) lo = mid; else hi = mid;
    }
    return lo;
}

// ---- nodes.  The reference numbers its nodes  nodes = [root] + desc(root),  desc(X) = [L, R] + desc(R) + desc(L)
// (bvh.cpp:156-161 pushes L then R, :115-116 pops the back), so for an internal node X whose children sit at l, l + 1:
//   desc(X) is the index range [l, end(X)),  end(root) = n_nodes,  end(L) = end(X),  end(R) = L internal ? L.left : end(X),
// and the subtree of X holds (end - left) / 2 internal nodes, X included.  The device numbers internal nodes in the order
// rtb_layout.hpp's convert_bvh does (a node's internal children get adjacent indices, the left subtree follows), which
// therefore follows from (left, end) alone, top down:  with f = the next free device index when X is reached,
//   dev(L) = f,  dev(R) = f + [L internal],  f(L) = f + [L int] + [R int],  f(R) = f(L) + (L int ? (end(L) - L.left) / 2 - 1 : 0).
// One frontier item per internal node; a level of the tree is one launch.
struct NodeItem {
    uint32_t ref, dev, next_free, end;
};

// sorted table of the leaves with more than 7 triangles (reference node index -> first entry of its run in big_leaves),
// built on the host in node-index order (rtb_layout.hpp: big_leaf_slots)
struct BigLeafIndex {
    const uint32_t *node;
    const uint32_t *slot;
    int n;
};

RTB_HD int leaf_link_of(const int32_t *l, uint32_t ref, uint32_t n_tris, const BigLeafIndex &big, int *err) {
    const int first = l[3], count = l[4];
    if (first < 0 || (uint64_t) first + (uint64_t) count > n_tris) { *err = kAsmLeafRange; return kLeafFlag | (1 << 28); }
    if (count <= 7) return kLeafFlag | (count << 28) | first;
    int lo = 0, hi = big.n - 1;
    while (lo <= hi) {
        const int mid = (lo + hi) >> 1;
        const uint32_t v = big.node[mid];
        if (v == ref) return kLeafFlag | (int) big.slot[mid];
        if (v < ref) lo = mid + 1; else hi = mid - 1;
    }
    *err = kAsmBigLeaf;
    return kLeafFlag | (1 << 28);
}

// Internal node `it` -> its 64-byte record; children that are internal are appended to out[] (out_n = how many: 0..2).
RTB_HD int convert_node(const NodeItem &it, uint32_t n_nodes, uint32_t n_internal, uint32_t n_tris, const float *aabb, const int32_t *links,
                        const BigLeafIndex &big, DNode *nodes, NodeItem out[2], int *out_n) {
    *out_n = 0;
    int err = kAsmOk;
    const int32_t *lx = links + (size_t) it.ref * 5;
    const int32_t l = lx[0];
    if (l < 0 || lx[1] != l + 1 || (uint32_t) l + 1u >= n_nodes || (uint32_t) l <= it.ref) return lx[1] < 0 || l < 0 ? kAsmChildren : kAsmNodeOrder;
    if (lx[2] < 0 || lx[2] > 2) return kAsmSplitDim;
    if (it.dev >= n_internal || it.end > n_nodes || it.end < (uint32_t) l + 2u) return kAsmNodeOrder;
    const uint32_t L = (uint32_t) l, R = L + 1u;
    const int32_t *ll = links + (size_t) L * 5, *lr = links + (size_t) R * 5;
    const bool Li = !(ll[4] > 0), Ri = !(lr[4] > 0);
    const uint32_t below = L + 2u;                                 // where desc(R) starts
    const uint32_t end_L = it.end, end_R = Li ? (uint32_t) ll[0] : it.end;
    // the order property, checked where it is used
    if (Ri && (uint32_t) lr[0] != below) err = kAsmNodeOrder;
    if (!Ri && Li && (uint32_t) ll[0] != below) err = kAsmNodeOrder;
    if (!Ri && !Li && it.end != below) err = kAsmNodeOrder;
    if (Li && (ll[0] < 0 || (uint32_t) ll[0] < below || (uint32_t) ll[0] + 2u > end_L || ((end_L - (uint32_t) ll[0]) & 1u))) err = kAsmNodeOrder;
    if (Ri && (end_R < below + 2u || ((end_R - below) & 1u))) err = kAsmNodeOrder;
    if (err != kAsmOk) return err;
    const uint32_t ni_L = Li ? (end_L - (uint32_t) ll[0]) / 2u - 1u : 0u;   // internal nodes strictly below L: |desc(L)| = 2 x (those + L itself)
    const uint32_t dev_L = it.next_free, dev_R = it.next_free + (Li ? 1u : 0u);
    const uint32_t f_L = it.next_free + (Li ? 1u : 0u) + (Ri ? 1u : 0u);
    const uint32_t f_R = f_L + ni_L;
    const int link_l = Li ? (int) dev_L : leaf_link_of(ll, L, n_tris, big, &err);
    const int link_r = Ri ? (int) dev_R : leaf_link_of(lr, R, n_tris, big, &err);
    const float *bl = aabb + (size_t) L * 6, *br = aabb + (size_t) R * 6;
    DNode nd;
    nd.a = make_float4(bl[0], bl[1], bl[2], bl[3]);
    nd.b = make_float4(bl[4], bl[5], br[0], br[1]);
    nd.c = make_float4(br[2], br[3], br[4], br[5]);
    nd.d = make_int4(link_l, link_r, lx[2], 0);
    nodes[it.dev] = nd;
    if (Li) out[(*out_n)++] = NodeItem{L, dev_L, f_L, end_L};
    if (Ri) out[(*out_n)++] = NodeItem{R, dev_R, f_R, end_R};
    return err;
}

// the bilinear footprint that starts at texel (x, y) of a row-major RGBA8 image, wrapped (DTexture::quads)
RTB_HD uint4 texture_quad(const uchar4 *texels, int width, int height, int x, int y) {
    const int x1 = x + 1 >= width ? x + 1 - width : x + 1, y1 = y + 1 >= height ? y + 1 - height : y + 1;
    const uint32_t *w = reinterpret_cast<const uint32_t *>(texels);
    return make_uint4(w[(size_t) y * width + x], w[(size_t) y1 * width + x], w[(size_t) y * width + x1], w[(size_t) y1 * width + x1]);
}

// texel (x, y) of a fused record (DFused): the quads of the three textures (null: the material has none there)
RTB_HD void fused_texel(const uchar4 *base, const uchar4 *mr, const uchar4 *nrm, int width, int height, int x, int y, uint4 out[4]) {
    const uint4 zero = make_uint4(0u, 0u, 0u, 0u);
    out[kFusedBase] = base ? texture_quad(base, width, height, x, y) : zero;
    out[kFusedMR] = mr ? texture_quad(mr, width, height, x, y) : zero;
    out[kFusedNormal] = nrm ? texture_quad(nrm, width, height, x, y) : zero;
    out[3] = zero;
}

inline const char *assemble_error_text(int e) {
    switch (e) {
        case kAsmVertexIndex: return "[gltf check] vertex index out of range";
        case kAsmSource: return "tri_source entry out of range";
        case kAsmNodeOrder: return "nodes are not in the reference's order";
        case kAsmLeafRange: return "leaf range out of bounds";
        case kAsmChildren: return "internal node without two children";
        case kAsmSplitDim: return "internal node with invalid split_dim";
        case kAsmBigLeaf: return "big-leaf table incomplete";
        default: return "unknown assembly error";
    }
}

}  // namespace rtb
