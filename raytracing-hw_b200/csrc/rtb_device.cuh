// Device-side data layout and the decision-exact traversal primitives (sm_100a).
//
// Arithmetic contract: the reference is built for the x86-64 baseline (SSE2, no FMA), so every float expression
// below is a restatement with the reference's operand order, and this translation unit is compiled with
// --fmad=false --prec-div=true --prec-sqrt=true --ftz=false.  Nothing here may be "simplified" algebraically.
//
//   ray/AABB : AABB::intersect       src/geometry/primitive.cpp:146-208  (Graphics Gems, returns entry distance)
//   ray/tri  : Primitive::intersect  src/geometry/primitive.cpp:17-57    (two-sided Moller-Trumbore)
//   closest  : BVH::intersectHelper  src/core/bvh.cpp:177-237            (near child by sign of direction[split_dim])
//   all hits : BVH::intersectAllHelper src/core/bvh.cpp:245-279
#pragma once
#include <cmath>
#include <cstdint>
#include <cstring>
#include <cuda_runtime.h>

// Every function below is usable from device code (the product: kernels in rtb_kernels.cu / rtb_integrator.cu) and,
// for CPU-side debugging of the restated arithmetic without a GPU, from host code compiled by the same nvcc
// (tests/devsim/ only -- librtb.so never calls them on the host; there is no CPU fallback behind the C ABI).
#define RTB_HD __host__ __device__ __forceinline__

namespace rtb {

template <class T>
RTB_HD T ldg(const T *p) {
#ifdef __CUDA_ARCH__
    return __ldg(p);
#else
    return *p;
#endif
}
RTB_HD float u2f(uint32_t u) {
#ifdef __CUDA_ARCH__
    return __uint_as_float(u);
#else
    float f;
    std::memcpy(&f, &u, 4);
    return f;
#endif
}
RTB_HD uint32_t f2u(float f) {
#ifdef __CUDA_ARCH__
    return __float_as_uint(f);
#else
    uint32_t u;
    std::memcpy(&u, &f, 4);
    return u;
#endif
}

// ---- 64-byte node: both child boxes + child links, one fetch (4 x LDG.128) serves the two box tests that
// bvh.cpp:185 and :193 do on nodes[left] and nodes[right].
//   a = lmin.x lmin.y lmin.z lmax.x | b = lmax.y lmax.z rmin.x rmin.y | c = rmin.z rmax.x rmax.y rmax.z
//   d = left link, right link, split axis, unused
// link >= 0: index of an internal node; link < 0: leaf, bits[30:28] = count (1..7), bits[27:0] = first triangle;
// count field 0 = entry `first` of the big-leaf table {first, count} (leaves with more than 7 triangles).
struct __align__(16) DNode {
    float4 a, b, c;
    int4 d;
};

constexpr int kLeafFlag = (int) 0x80000000u;
constexpr int kMaxStack = 64;
constexpr int kTriStride = 4;   // float4 per triangle record (the 4th is padding: a 48-byte record already costs two 32-byte sectors)
constexpr float kFar = 1e9f;   // `intersection.distance = 1e9` (bvh.cpp:181)

// A texture is read through `quads`: for every texel (x, y) the four taps of the bilinear footprint that starts there --
// {(x, y), (x, y + 1), (x + 1, y), (x + 1, y + 1)}, wrapped (Texture::interpolate_sample, primitive.h:182-215) -- as one
// 16-byte record, written on the device when the scene is created (rtb_assemble.cuh: texture_quad).  A bilinear fetch is
// then ONE 128-bit load = one 32-byte sector instead of four 4-byte loads in two rows (2.25 sectors on average); the shade
// step is bound by random 32-byte DRAM sectors.  4 x the bytes in HBM (1.1 GB for Sponza's 69
// textures), the same bytes over PCIe.  `texels` stays the row-major RGBA8 image the file delivered.
struct DTexture {
    const uchar4 *texels;
    const uint4 *quads;      // null when every use of the texture goes through a DFused record
    int width, height;
};
// The textures of ONE material that share a size -- base colour, metallic-roughness, normal map: all three are sampled at
// the same texture coordinate of every shaded hit (primitive.cpp:86-140) -- interleaved per texel: 64 bytes =
// {quad of the base colour, quad of metallic-roughness, quad of the normal map, unused}, so that the three bilinear
// fetches of a hit touch ONE 64-byte line (one DRAM burst) instead of three random ones.  Written on the device at scene
// creation (rtb_assemble.cu: k_fuse_textures).  16 x the RGBA8 bytes in HBM (1.7 GB for Sponza's 25 materials at 1024^2).
struct DFused {
    const uint4 *data;       // 4 x uint4 per texel: slot 0 base colour, 1 metallic-roughness, 2 normal map
    int width, height;
};
constexpr int kFusedBase = 0, kFusedMR = 1, kFusedNormal = 2;

struct DBvh {
    const DNode *nodes;       // internal nodes only
    const float4 *tris;       // kTriStride x float4 per triangle: (p0.xyz, U.x) (U.yz, V.xy) (V.z, Ng.xyz) (pad): 64-byte records,
                              // so that node and triangle records are both fetched with two 256-bit loads (rtb_trace.cuh)
    const int2 *big_leaves;   // {first, count}
    float root_box[6];
    int root_link;            // link of the root (a leaf link when the whole BVH is one leaf)
    int n_internal;
    int n_tris;
    int max_depth;
};

struct DScene {
    DBvh bvh;                 // scene BVH
    DBvh light_bvh;           // emissive triangles (unordered all-hits queries)
    int n_lights;
    const float *light_inv_area;      // 1.f / (0.5f * length(cross(U,V)))  per light triangle
    // cold shading data, per triangle (post-build order), 128-bit aligned records
    const float4 *tri_shade;  // 4 x float4 : n0.xyz n1.x | n1.yz n2.xy | n2.z uv0.xy uv1.x | uv1.y uv2.xy mesh_id(bits)
    const float4 *tri_tan;    // 3 x float4 : per-vertex tangents xyzw (null when no material has a normal map)
    const float4 *mesh_material;      // n_meshes x 4 float4 (16 floats, layout in rtb.h)
    const double *mesh_normal_transform;   // n_meshes x 16
    int n_meshes;
    const DTexture *textures;
    int n_textures;
    const DFused *fused;      // material record float 14 = index into this table, or -1
    float cam_pos[3];
    float cam_axes[9];
    float cam_tan_half_fov[2];
    int ray_depth;
    float max_distance;
    float bg_color[3];
    const float *lut;         // 512 floats in global memory (L1-resident gathers; a table in the kernel parameters would
                              // serialise divergent indices in the constant cache):
                              //   [0,256)   powf(i/255.f, 2.2f) evaluated by the HOST libm (primitive.h:208)
                              //   [256,512) (float) i / 255.f   (IEEE division, primitive.h:190-200)
};

struct Ray {
    float o[3], d[3], inv[3];
};

struct Hit {
    float t, u, v;
    int id;
};

// Ray::Ray (primitive.cpp:12-15): normalize(direction) = direction * (1.f / sqrt(((0+x*x)+y*y)+z*z)); inv = 1.f / d
RTB_HD void make_ray(Ray &r, float ox, float oy, float oz, float dx, float dy, float dz) {
    float mod = dx * dx;
    mod = mod + dy * dy;
    mod = mod + dz * dz;
    mod = 1.f / sqrtf(mod);
    r.o[0] = ox; r.o[1] = oy; r.o[2] = oz;
    r.d[0] = dx * mod; r.d[1] = dy * mod; r.d[2] = dz * mod;
    r.inv[0] = 1.f / r.d[0]; r.inv[1] = 1.f / r.d[1]; r.inv[2] = 1.f / r.d[2];
}
// same without inv_direction, for queries that are known not to run a ray/box test
RTB_HD void make_ray_no_inv(Ray &r, float ox, float oy, float oz, float dx, float dy, float dz) {
    float mod = dx * dx;
    mod = mod + dy * dy;
    mod = mod + dz * dz;
    mod = 1.f / sqrtf(mod);
    r.o[0] = ox; r.o[1] = oy; r.o[2] = oz;
    r.d[0] = dx * mod; r.d[1] = dy * mod; r.d[2] = dz * mod;
    r.inv[0] = 0.f; r.inv[1] = 0.f; r.inv[2] = 0.f;
}

// AABB::intersect (primitive.cpp:146-208).  Returns hit?; dist = entry distance (0 when the origin is inside).
RTB_HD bool box_test(const Ray &r, float mnx, float mny, float mnz, float mxx, float mxy, float mxz, float &dist) {
    const float mn[3] = {mnx, mny, mnz}, mx[3] = {mxx, mxy, mxz};
    float plane[3];
    bool mid[3];
    bool inside = true;
#pragma unroll
    for (int i = 0; i < 3; ++i) {
        const bool lo = r.o[i] < mn[i];
        const bool hi = r.o[i] > mx[i];
        plane[i] = lo ? mn[i] : mx[i];
        mid[i] = !(lo || hi);
        inside = inside && mid[i];
    }
    if (inside) {
        dist = 0.f;
        return true;
    }
    float maxT[3];
#pragma unroll
    for (int i = 0; i < 3; ++i)
        maxT[i] = (!mid[i] && r.d[i] != 0.f) ? (plane[i] - r.o[i]) * r.inv[i] : -1.f;
    // largest maxT, first maximum wins (strict <, scanning i = 1, 2)
    int which = 0;
    float tmax = maxT[0];
    if (tmax < maxT[1]) { which = 1; tmax = maxT[1]; }
    if (tmax < maxT[2]) { which = 2; tmax = maxT[2]; }
    if (tmax < 0.f) return false;
    float c[3];
    bool ok = true;
#pragma unroll
    for (int i = 0; i < 3; ++i) {
        const float ci = r.o[i] + tmax * r.d[i];
        const bool out = (ci < mn[i]) || (ci > mx[i]);
        if (which != i) {
            ok = ok && !out;
            c[i] = ci;
        } else {
            c[i] = plane[i];
        }
    }
    if (!ok) return false;
    // length(coord - origin): a - b == a + (-b) (vector.h:326-328), sum ((0+x*x)+y*y)+z*z
    const float ex = c[0] - r.o[0], ey = c[1] - r.o[1], ez = c[2] - r.o[2];
    float m = ex * ex;
    m = m + ey * ey;
    m = m + ez * ez;
    dist = sqrtf(m);
    return true;
}

// Primitive::intersect (primitive.cpp:17-57) on (p0, U, V); returns hit?, fills t,u,v.
RTB_HD bool tri_test(const Ray &r, const float4 q0, const float4 q1, const float4 q2, float &t, float &u, float &v) {
    const float p0x = q0.x, p0y = q0.y, p0z = q0.z;
    const float Ux = q0.w, Uy = q1.x, Uz = q1.y;
    const float Vx = q1.z, Vy = q1.w, Vz = q2.x;
    // cross(d, V) = (d1*V2 - V1*d2, d2*V0 - V2*d0, d0*V1 - V0*d1)   (vector.h:372-378)
    const float Px = r.d[1] * Vz - Vy * r.d[2];
    const float Py = r.d[2] * Vx - Vz * r.d[0];
    const float Pz = r.d[0] * Vy - Vx * r.d[1];
    float det = Ux * Px;
    det = det + Uy * Py;
    det = det + Uz * Pz;
    // `-1e-6 < det && det < 1e-6` with double literals <=> |det| <= float(1e-6) rejects
    if (fabsf(det) <= u2f(0x358637bdu)) return false;
    const float inv_det = 1.f / det;
    const float sx = r.o[0] - p0x, sy = r.o[1] - p0y, sz = r.o[2] - p0z;
    float su = sx * Px;
    su = su + sy * Py;
    su = su + sz * Pz;
    u = inv_det * su;
    if (u < 0.f || u > 1.f) return false;
    // cross(s, U)
    const float Qx = sy * Uz - Uy * sz;
    const float Qy = sz * Ux - Uz * sx;
    const float Qz = sx * Uy - Ux * sy;
    float sv = r.d[0] * Qx;
    sv = sv + r.d[1] * Qy;
    sv = sv + r.d[2] * Qz;
    v = inv_det * sv;
    if (v < 0.f || u + v > 1.f) return false;
    float st = Vx * Qx;
    st = st + Vy * Qy;
    st = st + Vz * Qz;
    t = inv_det * st;
    if (t < 0.f) return false;
    return true;
}

RTB_HD void decode_leaf(const DBvh &b, int link, int &first, int &count) {
    count = (link >> 28) & 7;
    first = link & 0x0FFFFFFF;
    if (count == 0) {
        const int2 bl = ldg(&b.big_leaves[first]);
        first = bl.x;
        count = bl.y;
    }
}

struct Counters {
    unsigned long long box, tri;
};

// BVH::intersect (bvh.cpp:239-243).  CULL = 0: the reference's sibling-only culling, reproduced exactly with a
// "min since push" value per deferred sibling; CULL = 1: global-best culling (same visit order, subset of visits).
template <int CULL, bool COUNT>
RTB_HD bool closest_hit_ordered(const DBvh &b, const Ray &r, Hit &best, Counters *cnt) {
    best.t = kFar;
    best.id = -1;
    best.u = 0.f;
    best.v = 0.f;
    float d0;
    if (COUNT) cnt->box += 1;
    if (!box_test(r, b.root_box[0], b.root_box[1], b.root_box[2], b.root_box[3], b.root_box[4], b.root_box[5], d0)) return false;

    int s_link[kMaxStack];
    float s_d[kMaxStack];
    float s_m[CULL == 0 ? kMaxStack : 1];
    int sp = 0;
    float cur = kFar, m_top = kFar;   // CULL==0 only
    int link = b.root_link;

    while (true) {
        if (link >= 0) {
            const DNode *np = b.nodes + link;
            const float4 na = ldg(&np->a), nb = ldg(&np->b), nc = ldg(&np->c);
            const int4 nd = ldg(&np->d);
            float dl, dr;
            const bool hl = box_test(r, na.x, na.y, na.z, na.w, nb.x, nb.y, dl);
            const bool hr = box_test(r, nb.z, nb.w, nc.x, nc.y, nc.z, nc.w, dr);
            if (COUNT) cnt->box += 2;
            const float dax = nd.z == 0 ? r.d[0] : (nd.z == 1 ? r.d[1] : r.d[2]);
            const bool left_first = dax > 0.f;   // bvh.cpp:183
            const int lA = left_first ? nd.x : nd.y, lB = left_first ? nd.y : nd.x;
            const float dA = left_first ? dl : dr, dB = left_first ? dr : dl;
            bool vA = left_first ? hl : hr, vB = left_first ? hr : hl;
            if (CULL == 1) {
                vA = vA && !(dA > best.t);
                vB = vB && !(dB > best.t);
            }
            if (vA) {
                if (vB) {
                    if (CULL == 0) {
                        if (sp > 0) s_m[sp - 1] = fminf(m_top, cur);
                        m_top = kFar;
                        cur = kFar;
                    }
                    s_link[sp] = lB;
                    s_d[sp] = dB;
                    ++sp;
                }
                link = lA;
                continue;
            }
            if (vB) {
                // near child missed: the far child is tested against a still-empty local best (1e9)
                if (CULL == 1 || !(dB > kFar)) {
                    link = lB;
                    continue;
                }
            }
        } else {
            int first, count;
            decode_leaf(b, link, first, count);
            for (int i = first; i < first + count; ++i) {
                const float4 *tp = b.tris + (size_t) i * kTriStride;
                const float4 q0 = ldg(tp), q1 = ldg(tp + 1), q2 = ldg(tp + 2);
                float t, u, v;
                if (COUNT) cnt->tri += 1;
                if (tri_test(r, q0, q1, q2, t, u, v)) {
                    if (CULL == 0) cur = fminf(cur, t);
                    if (t < best.t) {
                        best.t = t;
                        best.u = u;
                        best.v = v;
                        best.id = i;
                    }
                }
            }
        }
        // pop
        bool found = false;
        while (sp > 0) {
            --sp;
            link = s_link[sp];
            const float d = s_d[sp];
            float limit;
            if (CULL == 0) {
                limit = fminf(m_top, cur);     // best of the near sibling's subtree
                cur = limit;
                m_top = sp > 0 ? s_m[sp - 1] : kFar;
            } else {
                limit = best.t;
            }
            if (!(d > limit)) {   // `if (pre_inter.value() > intersection.distance) {} else recurse` (bvh.cpp:195-196)
                found = true;
                break;
            }
        }
        if (!found) break;
    }
    return best.id >= 0;
}

// BVH::intersectAll (bvh.cpp:245-279) fused with ManyLightsDistribution::pdf's accumulation
// (random.cpp:179-188): no root box test, left then right, no culling; visits hits in the reference's order so the
// float sum is identical.  `f(i, t)` is called for every triangle hit (t >= 0).
template <bool COUNT, class F>
RTB_HD void all_hits(const DBvh &b, const Ray &r, Counters *cnt, F f) {
    int s_link[kMaxStack];
    int sp = 0;
    int link = b.root_link;
    while (true) {
        if (link >= 0) {
            const DNode *np = b.nodes + link;
            const float4 na = ldg(&np->a), nb = ldg(&np->b), nc = ldg(&np->c);
            const int4 nd = ldg(&np->d);
            float dl, dr;
            const bool hl = box_test(r, na.x, na.y, na.z, na.w, nb.x, nb.y, dl);
            const bool hr = box_test(r, nb.z, nb.w, nc.x, nc.y, nc.z, nc.w, dr);
            if (COUNT) cnt->box += 2;
            if (hl) {
                if (hr) s_link[sp++] = nd.y;
                link = nd.x;
                continue;
            }
            if (hr) {
                link = nd.y;
                continue;
            }
        } else {
            int first, count;
            decode_leaf(b, link, first, count);
            for (int i = first; i < first + count; ++i) {
                const float4 *tp = b.tris + (size_t) i * kTriStride;
                const float4 q0 = ldg(tp), q1 = ldg(tp + 1), q2 = ldg(tp + 2);
                float t, u, v;
                if (COUNT) cnt->tri += 1;
                if (tri_test(r, q0, q1, q2, t, u, v)) f(i, t, q2);
            }
        }
        if (sp == 0) break;
        link = s_link[--sp];
    }
}

}  // namespace rtb
