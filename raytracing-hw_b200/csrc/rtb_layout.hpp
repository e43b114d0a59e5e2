// Host-side conversion of the flat scene description (include/rtb.h, reference order) into the device layout of
// rtb_device.cuh.  Plain host code shared by the uploader (rtb_api.cu) and the CPU-side debugging harness
// (tests/devsim/); nothing here touches CUDA memory.
#pragma once
#include "../../include/rtb.h"
#include "rtb_device.cuh"
#include "rtb_shade.cuh"

#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

namespace rtb {

struct HostBvhOut {
    std::vector<DNode> nodes;
    std::vector<float4> tris;
    std::vector<int2> big;
    std::vector<float> inv_area;
    float root_box[6];
    int root_link = 0;
    int max_depth = 0;
};

// Leaves of more than 7 triangles go through a table: a run of entries {first + 7k, count - 7k} per leaf (the link names
// the first one -- the whole leaf: what decode_leaf reads; the queue kernels take 7 triangles at a time and step along
// the run).  The runs are laid out in REFERENCE NODE ORDER, so that the host conversion below and the device-side one
// (rtb_assemble.cuh) produce the same table: big_node[k] = node index of the k-th such leaf (ascending), big_slot[k] =
// its first entry.
inline void big_leaf_slots(uint32_t n_nodes, const int32_t *links, uint32_t n_tris, std::vector<int2> &big, std::vector<uint32_t> &big_node,
                           std::vector<uint32_t> &big_slot) {
    big.clear(); big_node.clear(); big_slot.clear();
    for (uint32_t i = 0; i < n_nodes; ++i) {
        const int first = links[(size_t) i * 5 + 3], count = links[(size_t) i * 5 + 4];
        if (count <= 7 || (uint32_t) count > n_tris) continue;   // (a count beyond the array is rejected by the range check)
        big_node.push_back(i);
        big_slot.push_back((uint32_t) big.size());
        for (int k = 0; 7 * k < count; ++k) big.push_back(make_int2(first + 7 * k, count - 7 * k));
    }
}

struct LeafLinker {
    std::vector<uint32_t> big_node, big_slot;
    int link(uint32_t node, int first, int count) const {
        if (count >= 1 && count <= 7) return kLeafFlag | (count << 28) | first;
        const auto it = std::lower_bound(big_node.begin(), big_node.end(), node);
        if (it == big_node.end() || *it != node) return kLeafFlag | (1 << 28);   // (only for a leaf the range check rejects)
        return kLeafFlag | (int) big_slot[(size_t) (it - big_node.begin())];
    }
};

// reference-order nodes (aabb6, links5) -> device layout.  Internal nodes are numbered in depth-first (left first)
// order; triangle ids are untouched.
inline bool convert_bvh(uint32_t n_nodes, const float *aabb, const int32_t *links, uint32_t n_tris, const float *tri_pos, HostBvhOut &o,
                 std::string &err) {
    if (n_nodes == 0 || n_tris == 0) { err = "empty BVH"; return false; }
    if (n_tris >= (1u << 28)) { err = "too many triangles for the 28-bit leaf link"; return false; }
    // triangles: (p0.xyz, U.x) (U.yz, V.xy) (V.z, Ng.xyz); Ng = get_geometric_normal (primitive.cpp:77-84)
    o.tris.assign((size_t) n_tris * kTriStride, make_float4(0.f, 0.f, 0.f, 0.f));
    o.inv_area.resize(n_tris);
    for (uint32_t i = 0; i < n_tris; ++i) {
        const float *p = tri_pos + (size_t) i * 9;
        const float *U = p + 3, *V = p + 6;
        float n[3] = {U[1] * V[2] - V[1] * U[2], U[2] * V[0] - V[2] * U[0], U[0] * V[1] - V[0] * U[1]};
        float len = 0.f;
        for (int k = 0; k < 3; ++k) len += n[k] * n[k];
        len = std::sqrt(len);
        const float area = 0.5f * len;
        const float mod = 1.f / len;
        for (int k = 0; k < 3; ++k) n[k] *= mod;
        o.tris[(size_t) i * kTriStride + 0] = make_float4(p[0], p[1], p[2], U[0]);
        o.tris[(size_t) i * kTriStride + 1] = make_float4(U[1], U[2], V[0], V[1]);
        o.tris[(size_t) i * kTriStride + 2] = make_float4(V[2], n[0], n[1], n[2]);
        o.inv_area[i] = 1.f / area;   // `probability = 1.f / triangle_area` (random.cpp:88)
    }
    std::memcpy(o.root_box, aabb, 24);
    auto is_leaf = [&](uint32_t i) { return links[(size_t) i * 5 + 4] > 0; };
    auto check = [&](uint32_t i) -> bool {
        const int32_t *l = links + (size_t) i * 5;
        if (is_leaf(i)) {
            if (l[3] < 0 || (uint64_t) l[3] + (uint64_t) l[4] > n_tris) { err = "leaf range out of bounds"; return false; }
            return true;
        }
        if (l[0] < 0 || l[1] < 0 || (uint32_t) l[0] >= n_nodes || (uint32_t) l[1] >= n_nodes) { err = "internal node without two children"; return false; }
        if (l[2] < 0 || l[2] > 2) { err = "internal node with invalid split_dim"; return false; }
        return true;
    };
    if (!check(0)) return false;
    LeafLinker leaf;
    big_leaf_slots(n_nodes, links, n_tris, o.big, leaf.big_node, leaf.big_slot);
    if (is_leaf(0)) {
        o.root_link = leaf.link(0, links[3], links[4]);
        o.max_depth = 0;
        return true;
    }
    // number internal nodes depth-first
    struct Item { uint32_t ref; int dev; int depth; };
    std::vector<Item> stack;
    o.nodes.reserve(n_nodes / 2 + 1);
    o.nodes.emplace_back();
    stack.push_back({0u, 0, 1});
    o.root_link = 0;
    while (!stack.empty()) {
        const Item it = stack.back();
        stack.pop_back();
        o.max_depth = std::max(o.max_depth, it.depth);
        const int32_t *l = links + (size_t) it.ref * 5;
        const uint32_t L = (uint32_t) l[0], R = (uint32_t) l[1];
        if (!check(L) || !check(R)) return false;
        const float *bl = aabb + (size_t) L * 6, *br = aabb + (size_t) R * 6;
        int link_l, link_r;
        // right first onto the stack so that the left subtree gets the adjacent indices
        int dev_r = -1, dev_l = -1;
        if (!is_leaf(L)) { dev_l = (int) o.nodes.size(); o.nodes.emplace_back(); }
        if (!is_leaf(R)) { dev_r = (int) o.nodes.size(); o.nodes.emplace_back(); }
        link_l = is_leaf(L) ? leaf.link(L, links[(size_t) L * 5 + 3], links[(size_t) L * 5 + 4]) : dev_l;
        link_r = is_leaf(R) ? leaf.link(R, links[(size_t) R * 5 + 3], links[(size_t) R * 5 + 4]) : dev_r;
        if (dev_r >= 0) stack.push_back({R, dev_r, it.depth + 1});
        if (dev_l >= 0) stack.push_back({L, dev_l, it.depth + 1});
        DNode nd;
        nd.a = make_float4(bl[0], bl[1], bl[2], bl[3]);
        nd.b = make_float4(bl[4], bl[5], br[0], br[1]);
        nd.c = make_float4(br[2], br[3], br[4], br[5]);
        nd.d = make_int4(link_l, link_r, l[2], 0);
        o.nodes[(size_t) it.dev] = nd;
    }
    if (o.max_depth > kMaxStack) { err = "BVH deeper than the traversal stack (" + std::to_string(o.max_depth) + " > " + std::to_string(kMaxStack) + ")"; return false; }
    return true;
}

// cold shading records: 4 x float4 per triangle (see DScene::tri_shade)
inline void pack_shade(size_t n, const float *nrm9, const float *uv6, const int32_t *mesh, std::vector<float4> &out) {
    out.resize(n * 4);
    for (size_t i = 0; i < n; ++i) {
        float nn[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0}, uv[6] = {0, 0, 0, 0, 0, 0};
        if (nrm9) std::memcpy(nn, nrm9 + i * 9, 36);
        if (uv6) std::memcpy(uv, uv6 + i * 6, 24);
        const uint32_t m = mesh ? (uint32_t) mesh[i] : 0u;
        float mf;
        std::memcpy(&mf, &m, 4);
        out[i * 4 + 0] = make_float4(nn[0], nn[1], nn[2], nn[3]);
        out[i * 4 + 1] = make_float4(nn[4], nn[5], nn[6], nn[7]);
        out[i * 4 + 2] = make_float4(nn[8], uv[0], uv[1], uv[2]);
        out[i * 4 + 3] = make_float4(uv[3], uv[4], uv[5], mf);
    }
}

// Everything rtb_scene_create derives from a rtb_scene_desc, as host vectors in device layout.
struct HostLayout {
    HostBvhOut bvh, light_bvh;
    std::vector<float4> shade;              // DScene::tri_shade
    std::vector<float4> tan;                // DScene::tri_tan (empty when no material has a normal map)
    std::vector<float4> material;           // DScene::mesh_material
    std::vector<double> normal_transform;   // DScene::mesh_normal_transform
    int n_meshes = 0;
    bool any_normal_map = false;
    std::vector<float> lut;                 // DScene::lut
    // DScene::fused: the textures of a material that share a size, interleaved per texel (rtb_device.cuh: DFused)
    struct FusedGroup { int tex[3]; int width, height; };   // base colour, metallic-roughness, normal map (-1: none)
    std::vector<FusedGroup> fused;
    std::vector<char> texture_needs_quads;  // per texture: some use of it does not go through a fused record
};

// everything but the scene's own triangles and nodes: constants, light BVH (small, host-assembled), materials, textures
inline bool build_layout_common(const rtb_scene_desc *desc, HostLayout &L, std::string &err) {
    if (!desc->node_aabb || !desc->node_links || desc->n_tris == 0 || desc->n_nodes == 0) {
        err = "scene without nodes / triangles";
        return false;
    }
    if (desc->ray_depth < 0 || desc->ray_depth > kMaxRayDepth) {
        err = "ray_depth outside [0, " + std::to_string(kMaxRayDepth) + "]";
        return false;
    }
    if (desc->n_lights > 0) {
        if (!desc->light_pos || !desc->light_node_aabb || !desc->light_node_links || desc->n_light_nodes == 0) {
            err = "light list without its BVH";
            return false;
        }
        if (!convert_bvh(desc->n_light_nodes, desc->light_node_aabb, desc->light_node_links, desc->n_lights, desc->light_pos, L.light_bvh, err)) {
            err = "(lights) " + err;
            return false;
        }
    }
    std::vector<float> mat;
    if (desc->n_meshes > 0 && desc->mesh_material && desc->mesh_normal_transform) {
        mat.assign(desc->mesh_material, desc->mesh_material + (size_t) desc->n_meshes * 16);
        L.normal_transform.assign(desc->mesh_normal_transform, desc->mesh_normal_transform + (size_t) desc->n_meshes * 16);
    } else {   // struct Material defaults (primitive.h:76-89)
        const float m0[16] = {1.f, 1.f, 1.f, 1.f, 1.f, 0.f, 0.f, 0.f, 1.f, 1.f, -1.f, -1.f, -1.f, -1.f, 0.f, 0.f};
        mat.assign(m0, m0 + 16);
        const double I[16] = {1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1};
        L.normal_transform.assign(I, I + 16);
    }
    L.n_meshes = (int) (mat.size() / 16);
    L.any_normal_map = false;
    for (int m = 0; m < L.n_meshes; ++m) {
        for (int k = 10; k < 14; ++k) {
            const int ti = (int) mat[(size_t) m * 16 + k];
            if (ti >= (int) desc->n_textures) { err = "texture index out of range"; return false; }
            if (ti < 0) mat[(size_t) m * 16 + k] = -1.f;
        }
        L.any_normal_map = L.any_normal_map || mat[(size_t) m * 16 + 11] >= 0.f;
    }
    for (uint32_t i = 0; i < desc->n_textures; ++i) {
        const rtb_texture_desc &t = desc->textures[i];
        if (t.width <= 0 || t.height <= 0 || !t.rgba) { err = "bad texture"; return false; }
    }
    // fused records: a material with at least two of {base colour, metallic-roughness, normal map}, all of one size;
    // materials with the same three textures share a record (RTB_FUSE_TEXTURES=0: none -- every texture keeps its own quads)
    L.fused.clear();
    L.texture_needs_quads.assign(desc->n_textures, 0);
    static const bool fuse_on = [] { const char *e = std::getenv("RTB_FUSE_TEXTURES"); return !(e && std::atoi(e) == 0); }();
    for (int m = 0; m < L.n_meshes; ++m) {
        float *mm = &mat[(size_t) m * 16];
        const int t3[3] = {(int) mm[10], (int) mm[12], (int) mm[11]};   // base, metallic-roughness, normal
        int present = 0, w = 0, h = 0;
        bool same = true;
        for (int k = 0; k < 3; ++k)
            if (t3[k] >= 0) {
                const rtb_texture_desc &t = desc->textures[t3[k]];
                if (present && (t.width != w || t.height != h)) same = false;
                w = t.width;
                h = t.height;
                ++present;
            }
        int id = -1;
        if (fuse_on && present >= 2 && same) {
            for (size_t g = 0; g < L.fused.size(); ++g)
                if (L.fused[g].tex[0] == t3[0] && L.fused[g].tex[1] == t3[1] && L.fused[g].tex[2] == t3[2]) id = (int) g;
            if (id < 0) {
                id = (int) L.fused.size();
                L.fused.push_back({{t3[0], t3[1], t3[2]}, w, h});
            }
        } else {
            for (int k = 0; k < 3; ++k)
                if (t3[k] >= 0) L.texture_needs_quads[(size_t) t3[k]] = 1;
        }
        if ((int) mm[13] >= 0) L.texture_needs_quads[(size_t) (int) mm[13]] = 1;   // emission
        mm[14] = (float) id;
    }
    L.material.resize((size_t) L.n_meshes * 4);
    std::memcpy(L.material.data(), mat.data(), mat.size() * sizeof(float));
    return true;
}

inline bool build_layout(const rtb_scene_desc *desc, HostLayout &L, std::string &err) {
    if (!desc->tri_pos) { err = "scene without nodes / triangles"; return false; }
    if (!build_layout_common(desc, L, err)) return false;
    if (!convert_bvh(desc->n_nodes, desc->node_aabb, desc->node_links, desc->n_tris, desc->tri_pos, L.bvh, err)) return false;
    const size_t nt = desc->n_tris;
    if (desc->tri_mesh)
        for (size_t i = 0; i < nt; ++i)
            if (desc->tri_mesh[i] < 0 || desc->tri_mesh[i] >= L.n_meshes) { err = "mesh id out of range"; return false; }
    pack_shade(nt, desc->tri_nrm, desc->tri_uv, desc->tri_mesh, L.shade);
    if (L.any_normal_map) {
        if (!desc->tri_tan) { err = "normal-mapped material without tangents"; return false; }
        L.tan.resize(nt * 3);
        std::memcpy(L.tan.data(), desc->tri_tan, nt * 12 * sizeof(float));
    }
    return true;
}

// camera, integrator constants and the sRGB decode table (powf of the HOST libm, as the reference's
// Texture::interpolate_sample evaluates it per tap, primitive.h:208)
inline void fill_scalars(const rtb_scene_desc *desc, DScene &d) {
    d.n_lights = (int) desc->n_lights;
    d.n_textures = (int) desc->n_textures;
    std::memcpy(d.cam_pos, desc->cam_pos, 12);
    std::memcpy(d.cam_axes, desc->cam_axes, 36);
    d.cam_tan_half_fov[0] = desc->cam_tan_half_fov[0];
    d.cam_tan_half_fov[1] = desc->cam_tan_half_fov[1];
    d.ray_depth = desc->ray_depth;
    d.max_distance = desc->max_distance;
    std::memcpy(d.bg_color, desc->bg_color, 12);
}
inline void fill_lut(std::vector<float> &lut) {
    lut.resize(512);
    for (int i = 0; i < 256; ++i) {
        lut[(size_t) i] = std::pow((float) i / 255.f, 2.2f);
        lut[(size_t) 256 + i] = (float) i / 255.f;
    }
}
inline void fill_bvh_scalars(const HostBvhOut &h, DBvh &d, uint32_t n_tris) {
    std::memcpy(d.root_box, h.root_box, 24);
    d.root_link = h.root_link;
    d.n_internal = (int) h.nodes.size();
    d.n_tris = (int) n_tris;
    d.max_depth = h.max_depth;
}

}  // namespace rtb
