// Sample stream, sampling strategies, materials and the per-sample bounce loop, restated for device code.
//
// Decision-exact contract (see rtb_device.cuh): built with --fmad=false; every expression keeps the operand order
// and the float/double types the reference's x86-64 (SSE2, no FMA) build evaluates, so that for the same per-pixel
// engine seed a path takes the same decisions and produces the same radiance bits as the reference's
// Scene::intersect.  Only glibc's logf sits between the engine and a Gaussian draw; it is restated in rtb_libm.cuh.
//
//   engine / distributions : std::minstd_rand, uniform_real_distribution<float>, normal_distribution<float>
//                            (libstdc++ bits/random.h, random.tcc:1812-1845,3349-3380; src/utils/random.h:12-17)
//   sampling + pdfs        : src/utils/random.cpp:28-31,48-59,65-91,103-154,169-188,194-218
//   materials / textures   : src/geometry/primitive.cpp:86-140, src/geometry/primitive.h:182-215
//   BRDF                   : src/utils/vector.h:380-393,433-471, src/core/scene.cpp:134-154
//   bounce loop            : src/core/scene.cpp:71-156 (recursive there, an explicit per-bounce stack here)
//   camera / resolve       : src/core/camera.cpp:49-62, src/core/scene.cpp:54-63, src/utils/vector.h:222-233,400-407
#pragma once
#include "rtb_device.cuh"
#include "rtb_libm.cuh"
#include "rtb_traverse.cuh"

namespace rtb {

constexpr int kMaxRayDepth = 8;          // the parser fixes ray_depth = 6 (scene_parser.cpp:17)
constexpr float kStep = 1e-4f;           // `const float step = 1e-4;` (scene.cpp:11)
constexpr float kRoughness2Limit = 0.03f;   // scene.cpp:15
constexpr float kPi = 3.14159265358979323846264338327950288f;       // M_PIf32
constexpr float kInvPi = 0.318309886183790671537767526745028724f;   // M_1_PIf32

struct V3 {
    float x, y, z;
};
RTB_HD V3 v3(float x, float y, float z) { return V3{x, y, z}; }
RTB_HD V3 operator+(V3 a, V3 b) { return V3{a.x + b.x, a.y + b.y, a.z + b.z}; }
RTB_HD V3 operator-(V3 a) { return V3{-a.x, -a.y, -a.z}; }
RTB_HD V3 operator-(V3 a, V3 b) { return a + (-b); }                       // vector.h:326-328
RTB_HD V3 operator*(V3 v, float t) { return V3{v.x * t, v.y * t, v.z * t}; }   // vector.h:330-334
RTB_HD V3 operator*(float t, V3 v) { return v * t; }                        // vector.h:336-338
RTB_HD V3 operator*(V3 a, V3 b) { return V3{a.x * b.x, a.y * b.y, a.z * b.z}; }
RTB_HD float dot(V3 a, V3 b) {            // vector.h:365-370: s = 0; s += a[i]*b[i]
    float s = 0.f;
    s = s + a.x * b.x;
    s = s + a.y * b.y;
    s = s + a.z * b.z;
    return s;
}
RTB_HD V3 cross(V3 a, V3 b) {             // vector.h:372-378
    return V3{a.y * b.z - b.y * a.z, a.z * b.x - b.z * a.x, a.x * b.y - b.x * a.y};
}
// The shade kernel is 72 KB of SASS of which 44 KB are hot, against a 32 KB L1.5 instruction cache: 21 % of its stall samples are
// "no instruction" (profiles/r02h_sponza_k_wf_shade_summary.txt).  The helpers that were inlined at many sites are therefore
// can be kept as ONE copy each in device code with -DRTB_SHADE_NOINLINE=1.  Measured on one box, A/B (profiles/
// r02j_shade_noinline_ab.txt): the kernel shrinks 72 -> 66 KB, the calls cost more than the fetches they save -- shade
// step 134 -> 144 ms per 64-spp frame.  Off.
#ifndef RTB_SHADE_NOINLINE
#define RTB_SHADE_NOINLINE 0
#endif
#if defined(__CUDA_ARCH__) && RTB_SHADE_NOINLINE
#define RTB_HD_BIG __host__ __device__ __noinline__
#else
#define RTB_HD_BIG RTB_HD
#endif
RTB_HD float length(V3 v) {
    float m = 0.f;
    m = m + v.x * v.x;
    m = m + v.y * v.y;
    m = m + v.z * v.z;
    return sqrtf(m);
}
RTB_HD_BIG V3 normal(V3 v) {                  // vector.h:169-174
    float mod = length(v);
    mod = 1.f / mod;
    return V3{v.x * mod, v.y * mod, v.z * mod};
}
RTB_HD float std_max(float a, float b) { return (a < b) ? b : a; }   // std::max(a, b)
RTB_HD float std_min(float a, float b) { return (b < a) ? b : a; }   // std::min(a, b)

struct Q4 {
    float x, y, z, w;
};
// quat_from_two_vectors (vector.h:454-458) -> normal(vector4f)
RTB_HD Q4 quat_from_two_vectors(V3 u, V3 v) {
    const V3 w = cross(u, v);
    const float qw = 1.f + dot(u, v);
    float m = 0.f;
    m = m + w.x * w.x;
    m = m + w.y * w.y;
    m = m + w.z * w.z;
    m = m + qw * qw;
    float mod = sqrtf(m);
    mod = 1.f / mod;
    return Q4{w.x * mod, w.y * mod, w.z * mod, qw * mod};
}
// rotate(v, q) (vector.h:380-386): 2.f*dot(u,v)*u + (s*s - dot(u,u))*v + 2.f*s*cross(u,v)
RTB_HD V3 rotate(V3 v, Q4 q) {
    const V3 u = V3{q.x, q.y, q.z};
    const float s = q.w;
    return ((2.f * dot(u, v)) * u + (s * s - dot(u, u)) * v) + (2.f * s) * cross(u, v);
}

// ------------------------------------------------------------------------------------------------ sample stream
struct Rng {
    uint32_t x;          // minstd_rand state
    uint32_t has_saved;  // normal_distribution::_M_saved_available
    float saved;         // normal_distribution::_M_saved
};
RTB_HD void rng_seed(Rng &g, uint64_t seed) {   // linear_congruential_engine::seed: s mod m, 0 -> 1
    uint32_t s = (uint32_t) (seed % 2147483647ull);
    g.x = s == 0 ? 1u : s;
    g.has_saved = 0;
    g.saved = 0.f;
}
RTB_HD uint32_t rng_next(Rng &g) {   // x <- 48271 x mod (2^31 - 1)
    const uint64_t p = (uint64_t) g.x * 48271ull;
    uint32_t r = (uint32_t) (p & 0x7fffffffu) + (uint32_t) (p >> 31);
    if (r >= 0x7fffffffu) r -= 0x7fffffffu;
    g.x = r;
    return r;
}
// generate_canonical<float,24>: float(x - min) / 2147483648.f, clamped below 1 (random.tcc:3349-3380)
RTB_HD float rng_canonical(Rng &g) {
    const uint32_t v = rng_next(g) - 1u;
#ifdef __CUDA_ARCH__
    float u = __uint2float_rn(v) * 4.656612873077392578125e-10f;   // exact: division by 2^31
#else
    float u = (float) v * 4.656612873077392578125e-10f;
#endif
    if (u >= 1.f) u = u2f(0x3f7fffffu);
    return u;
}
RTB_HD float rng_uniform_pm1(Rng &g) { return rng_canonical(g) * 2.f + -1.f; }     // uniDist(-1,1) (random.cpp:21)
RTB_HD float rng_pixel_offset(Rng &g) { return rng_canonical(g) * 1.f + -0.5f; }   // offset(-0.5,0.5) (scene.cpp:18)
// normal_distribution<float>(0,1)::operator() (random.tcc:1812-1845), Marsaglia polar with a cached second value
RTB_HD float rng_normal(Rng &g) {
    float ret;
    if (g.has_saved) {
        g.has_saved = 0;
        ret = g.saved;
    } else {
        float x, y, r2;
        do {
            x = 2.f * rng_canonical(g) - 1.f;
            y = 2.f * rng_canonical(g) - 1.f;
            r2 = x * x + y * y;
        } while (r2 > 1.f || r2 == 0.f);
        const float mult = sqrtf(-2.f * glibc_logf(r2) / r2);
        g.saved = x * mult;
        g.has_saved = 1;
        ret = y * mult;
    }
    return ret * 1.f + 0.f;
}

// ------------------------------------------------------------------------------------------------ sampling
// uniform::sphere (random.cpp:28-31)
RTB_HD V3 sample_sphere(Rng &g) {
    V3 a;
    a.x = rng_normal(g);
    a.y = rng_normal(g);
    a.z = rng_normal(g);
    return normal(a);
}
// cosine_weighted::sample (random.cpp:48-55); `length(dir) < 1e-12` in double <=> len <= 0x2b8cbccc as float
RTB_HD V3 sample_cosine(V3 n, Rng &g) {
    V3 dir;
    do {
        dir = sample_sphere(g) + n;
    } while (length(dir) <= u2f(0x2b8cbcccu));
    float mod = length(dir);
    mod = 1.f / mod;
    dir.x = dir.x * mod;
    dir.y = dir.y * mod;
    dir.z = dir.z * mod;
    return dir;
}
RTB_HD float pdf_cosine(V3 n, V3 dir) { return std_max(0.f, dot(dir, n)) * kInvPi; }   // random.cpp:57-59

// visible_normal::sample (random.cpp:103-134)
RTB_HD V3 sample_vndf(V3 n, V3 eye, float alpha, Rng &g) {
    const V3 Z = V3{0.f, 0.f, 1.f};
    const Q4 rot = quat_from_two_vectors(n, Z);
    const V3 Ve = rotate(eye, rot);
    const V3 Vh = normal(V3{Ve.x * alpha, Ve.y * alpha, Ve.z});
    const float lensq = Vh.x * Vh.x + Vh.y * Vh.y;
    const V3 T1 = lensq > 0.f ? V3{-Vh.y, Vh.x, 0.f} * (1.f / sqrtf(lensq)) : V3{1.f, 0.f, 0.f};
    const V3 T2 = cross(Vh, T1);
    float tx, ty;
    do {
        tx = rng_uniform_pm1(g);
        ty = rng_uniform_pm1(g);
    } while (tx * tx + ty * ty > 1.f);
    const float s = 0.5f + 0.5f * Vh.z;
    ty = (1.f - s) * sqrtf(1.f - tx * tx) + s * ty;
    const V3 Nh = (tx * T1 + ty * T2) + sqrtf(std_max(0.f, 1.f - tx * tx - ty * ty)) * Vh;
    V3 Ne = normal(V3{alpha * Nh.x, alpha * Nh.y, std_max(0.f, Nh.z)});
    Ne = rotate(Ne, Q4{-rot.x, -rot.y, -rot.z, rot.w});
    return normal((2.f * Ne) * dot(Ne, eye) - eye);
}
// visible_normal::pdf (random.cpp:136-154)
RTB_HD float pdf_vndf(V3 n, V3 eye, float alpha, V3 dir) {
    const V3 Z = V3{0.f, 0.f, 1.f};
    const Q4 rot = quat_from_two_vectors(n, Z);
    const V3 sn = normal(dir + eye);
    const V3 V = rotate(eye, rot);
    const V3 Ni = rotate(sn, rot);
    if (dot(V, Ni) < 0.f) return 0.f;
    const float alpha2 = alpha * alpha;
    // M_PIf32 * alpha * alpha * std::pow(<float>, 2): the pow is the double overload, x*x is exact in double
    const float arg = Ni.x * Ni.x / alpha2 + Ni.y * Ni.y / alpha2 + Ni.z * Ni.z;
    const float invD = (float) ((double) (kPi * alpha * alpha) * ((double) arg * (double) arg));
    const float invG1 = 0.5f + 0.5f * sqrtf(1.f + (alpha2 * V.x * V.x + alpha2 * V.y * V.y) / (V.z * V.z));
    return 1.f / (4.f * invD * invG1 * dot(V, Z));
}

// ManyLightsDistribution::sample -> LightDistribution::sample (random.cpp:169-176, 65-81)
RTB_HD V3 sample_light(const DScene &sc, V3 point, Rng &g) {
    int pick = (int) floorf((rng_uniform_pm1(g) + 1.f) * 0.5f * (float) sc.n_lights);
    if (pick == sc.n_lights) pick -= 1;
    float u = (rng_uniform_pm1(g) + 1.f) / 2.f;
    float v = (rng_uniform_pm1(g) + 1.f) / 2.f;
    if (u + v > 1.f) {
        u = 1.f - u;
        v = 1.f - v;
    }
    const float4 *tp = sc.light_bvh.tris + (size_t) pick * kTriStride;
    const float4 q0 = ldg(tp), q1 = ldg(tp + 1), q2 = ldg(tp + 2);
    const V3 o = V3{q0.x, q0.y, q0.z}, U = V3{q0.w, q1.x, q1.y}, Vv = V3{q1.z, q1.w, q2.x};
    const V3 res = (o + u * U) + v * Vv;
    return normal(res - point);
}
// ManyLightsDistribution::pdf (random.cpp:179-188): sum over ALL light triangles the (unoccluded) ray hits
template <bool COUNT>
RTB_HD float pdf_lights(const DScene &sc, V3 point, V3 dir, Counters *cnt) {
    Ray r;
    // Ray{point, direction} re-normalises; inv_direction only feeds ray/box tests, and a light BVH that is a single
    // leaf (two light triangles in Sponza, practice7_*) runs none (intersectAll does not test the root box)
    if (sc.light_bvh.root_link >= 0) make_ray(r, point.x, point.y, point.z, dir.x, dir.y, dir.z);
    else make_ray_no_inv(r, point.x, point.y, point.z, dir.x, dir.y, dir.z);
    float prob = 0.f;
    const float *inv_area = sc.light_inv_area;
    all_hits<COUNT>(sc.light_bvh, r, cnt, [&](int li, float t, const float4 &q2) {
        // pdfWithHint (random.cpp:83-91): abs(probability * (t*t) / dot(normal, direction)); abs absorbs the flip
        const float dn = dot(V3{q2.y, q2.z, q2.w}, dir);
        prob = prob + fabsf(ldg(&inv_area[li]) * (t * t) / dn);
    });
    return prob / (float) sc.n_lights;
}

// SceneDistribution::sample (random.cpp:194-208)
RTB_HD V3 sample_direction(const DScene &sc, V3 point, V3 n, V3 eye, float alpha, Rng &g) {
    float s = (rng_uniform_pm1(g) + 1.f) * 3.f * 0.5f;
    if (sc.n_lights == 0) s = s / 1.5f;
    if (s <= 1.f) return sample_cosine(n, g);
    if (s <= 2.f) return sample_vndf(n, eye, alpha, g);
    return sample_light(sc, point, g);
}
// SceneDistribution::pdf (random.cpp:210-218)
template <bool COUNT>
RTB_HD float pdf_direction(const DScene &sc, V3 point, V3 n, V3 eye, float alpha, V3 dir, Counters *cnt, unsigned long long *light_queries) {
    if (sc.n_lights == 0) return (pdf_cosine(n, dir) + pdf_vndf(n, eye, alpha, dir)) / 2.f;
    const float pc = pdf_cosine(n, dir);
    *light_queries += 1;
    const float pl = pdf_lights<COUNT>(sc, point, dir, cnt);
    const float pv = pdf_vndf(n, eye, alpha, dir);
    return ((pc + pl) + pv) / 3.f;
}

// ------------------------------------------------------------------------------------------------ textures
// Texture::interpolate_sample (primitive.h:182-215), channels = 4 (tinygltf always expands), repeat wrap.
// SRGB: per-tap pow(v, 2.2f) through the host-computed 256-entry table (same glibc powf as the reference).
RTB_HD int wrap_index(int i, int n) {   // i % n for 0 <= i <= n + 1
    i = i >= n ? i - n : i;
    return i >= n ? i - n : i;
}
RTB_HD uchar4 u2c(uint32_t v) { return make_uchar4((unsigned char) (v & 255u), (unsigned char) ((v >> 8) & 255u), (unsigned char) ((v >> 16) & 255u), (unsigned char) (v >> 24)); }
// where the quads of a texture use live: the material's fused record (slot k of 4 per texel) or the texture's own array
struct TexView {
    const uint4 *q;
    int stride;       // uint4 per texel
    int width, height;
    const float *lut; // DScene::lut
};
RTB_HD TexView tex_view(const DScene &sc, int tex, int fused, int slot) {
    if (fused >= 0) {
        const DFused f = sc.fused[fused];
        return TexView{f.data + slot, 4, f.width, f.height, sc.lut};
    }
    const DTexture t = sc.textures[tex];
    return TexView{t.quads, 1, t.width, t.height, sc.lut};
}
template <bool SRGB>
RTB_HD_BIG void texture_sample(const TexView &t, float px, float py, float out[4]) {
    px = px - floorf(px);
    py = py - floorf(py);
    px = px * (float) t.width;
    py = py * (float) t.height;
    const int ix = (int) floorf(px), iy = (int) floorf(py);
    const float dx = px - floorf(px), dy = py - floorf(py);
    // 0 <= ix <= width (px lies in [0,1] after the wrap), so `% width` is at most two conditional subtractions; the taps at
    // (x0 + 1) % width and (y0 + 1) % height ride in the quad of (x0, y0)
    const int x0 = wrap_index(ix, t.width);
    const int y0 = wrap_index(iy, t.height);
    const uint4 q = ldg(&t.q[((size_t) y0 * t.width + x0) * (size_t) t.stride]);
    const uchar4 c00 = u2c(q.x);   // (dx=0, dy=0)
    const uchar4 c01 = u2c(q.y);   // (dx=0, dy=1)
    const uchar4 c10 = u2c(q.z);   // (dx=1, dy=0)
    const uchar4 c11 = u2c(q.w);   // (dx=1, dy=1)
    const unsigned char b00[4] = {c00.x, c00.y, c00.z, c00.w}, b01[4] = {c01.x, c01.y, c01.z, c01.w};
    const unsigned char b10[4] = {c10.x, c10.y, c10.z, c10.w}, b11[4] = {c11.x, c11.y, c11.z, c11.w};
#pragma unroll
    for (int i = 0; i < 3; ++i) {   // the alpha channel is never read by the hot path (.reduce())
        float a00, a01, a10, a11;
        if (SRGB) {
            a00 = ldg(&t.lut[b00[i]]); a01 = ldg(&t.lut[b01[i]]); a10 = ldg(&t.lut[b10[i]]); a11 = ldg(&t.lut[b11[i]]);
        } else {   // (float) byte / 255.f, tabulated
            a00 = ldg(&t.lut[256 + b00[i]]); a01 = ldg(&t.lut[256 + b01[i]]); a10 = ldg(&t.lut[256 + b10[i]]); a11 = ldg(&t.lut[256 + b11[i]]);
        }
        // interpolate_bilinear (primitive.h:158-161)
        out[i] = a00 * (1.f - dx) * (1.f - dy) + a01 * (1.f - dx) * dy + a10 * dx * (1.f - dy) + a11 * dx * dy;
    }
    out[3] = 0.f;
}

struct Material {   // struct Material (primitive.h:76-89) as uploaded: 16 floats
    float ior, alpha;
    V3 base, emission;
    float metallic, roughness2;
    int base_i, normal_i, mr_i, emission_i;
    int fused;      // DScene::fused entry that holds base / metallic-roughness / normal map, or -1
};
RTB_HD Material load_material(const DScene &sc, int mesh) {
    const float4 *mp = sc.mesh_material + (size_t) mesh * 4;
    const float4 a = ldg(mp), b = ldg(mp + 1), c = ldg(mp + 2), d = ldg(mp + 3);
    Material m;
    m.ior = a.x; m.alpha = a.y;
    m.base = V3{a.z, a.w, b.x};
    m.emission = V3{b.y, b.z, b.w};
    m.metallic = c.x; m.roughness2 = c.y;
    m.base_i = (int) c.z; m.normal_i = (int) c.w; m.mr_i = (int) d.x; m.emission_i = (int) d.y;
    m.fused = (int) d.z;
    return m;
}

struct Shade {      // the cold per-triangle record, unpacked
    V3 n0, n1, n2;
    float uv[6];
    int mesh;
};
RTB_HD Shade load_shade(const DScene &sc, int tri) {
    const float4 *p = sc.tri_shade + (size_t) tri * 4;
    const float4 a = ldg(p), b = ldg(p + 1), c = ldg(p + 2), d = ldg(p + 3);
    Shade s;
    s.n0 = V3{a.x, a.y, a.z};
    s.n1 = V3{a.w, b.x, b.y};
    s.n2 = V3{b.z, b.w, c.x};
    s.uv[0] = c.y; s.uv[1] = c.z; s.uv[2] = c.w; s.uv[3] = d.x; s.uv[4] = d.y; s.uv[5] = d.z;
    s.mesh = (int) f2u(d.w);
    return s;
}
// Primitive::get_texcoord (primitive.cpp:107-109): (1-u-v)*tc0 + u*tc1 + v*tc2 on vector2f
RTB_HD void texcoord(const Shade &s, float u, float v, float &tx, float &ty) {
    const float w = 1.f - u - v;
    tx = (s.uv[0] * w + s.uv[2] * u) + s.uv[4] * v;
    ty = (s.uv[1] * w + s.uv[3] * u) + s.uv[5] * v;
}

// multiplyVector(matrix4d, vector3f) (matrix.h:66-86): float accumulators, double products, t.w = 0
RTB_HD V3 multiply_vector(const double *m, V3 t) {
    const float tv[4] = {t.x, t.y, t.z, 0.f};
    float res[3];
#pragma unroll
    for (int i = 0; i < 3; ++i) {
        float acc = 0.f;
#pragma unroll
        for (int j = 0; j < 4; ++j) acc = (float) ((double) acc + ldg(&m[4 * j + i]) * (double) tv[j]);
        res[i] = acc;
    }
    return V3{res[0], res[1], res[2]};
}

// Primitive::get_shading_normal (primitive.cpp:86-105)
RTB_HD V3 shading_normal(const DScene &sc, int tri, const Shade &s, const Material &m, float u, float v) {
    const float w = 1.f - u - v;
    const V3 local_z = normal((w * s.n0 + u * s.n1) + v * s.n2);
    if (m.normal_i < 0) return local_z;
    const float4 *tp = sc.tri_tan + (size_t) tri * 3;
    const float4 t0 = ldg(tp), t1 = ldg(tp + 1), t2 = ldg(tp + 2);
    const V3 tin = normal((w * V3{t0.x, t0.y, t0.z} + u * V3{t1.x, t1.y, t1.z}) + v * V3{t2.x, t2.y, t2.z});
    const V3 local_x = normal(multiply_vector(sc.mesh_normal_transform + (size_t) s.mesh * 16, tin));
    const V3 local_y = cross(local_z, local_x) * t0.w;
    float tx, ty;
    texcoord(s, u, v, tx, ty);
    float smp[4];
    texture_sample<false>(tex_view(sc, m.normal_i, m.fused, kFusedNormal), tx, ty, smp);
    const V3 ln = V3{smp[0] + -0.5f, smp[1] + -0.5f, smp[2] + -0.5f} * 2.f;
    return normal((local_x * ln.x + local_y * ln.y) + local_z * ln.z);
}
// Primitive::get_color / get_emission (primitive.cpp:111-129)
// fused >= 0: the texture is slot kFusedBase of that fused record (the base colour); -1: its own quads (emission)
RTB_HD V3 textured_factor(const DScene &sc, const Shade &s, int tex, int fused, V3 factor, float u, float v) {
    if (tex < 0) return factor;
    float tx, ty;
    texcoord(s, u, v, tx, ty);
    float smp[4];
    texture_sample<true>(tex_view(sc, tex, fused, kFusedBase), tx, ty, smp);
    return V3{smp[0], smp[1], smp[2]} * factor;
}
// Primitive::get_metallic_roughness (primitive.cpp:131-140) -> {roughness2, metallic}
RTB_HD void metallic_roughness(const DScene &sc, const Shade &s, const Material &m, float u, float v, float &rough2, float &metallic) {
    if (m.mr_i < 0) {
        rough2 = m.roughness2;
        metallic = m.metallic;
        return;
    }
    float tx, ty;
    texcoord(s, u, v, tx, ty);
    float smp[4];
    texture_sample<false>(tex_view(sc, m.mr_i, m.fused, kFusedMR), tx, ty, smp);
    const float r2 = smp[1] * smp[1];
    rough2 = r2 * m.roughness2;
    metallic = smp[2] * m.metallic;
}

// ------------------------------------------------------------------------------------------------ BRDF
RTB_HD float ggx_d(float alpha, V3 N, V3 H) {   // vector.h:433-439
    const float alpha2 = alpha * alpha;
    const float NdotH = dot(N, H);
    const float x = NdotH * NdotH * (alpha2 - 1.f) + 1.f;
    return alpha2 * kInvPi / (x * x);
}
RTB_HD float smith_g(float alpha, V3 N, V3 V, V3 L) {   // vector.h:442-452
    const float alpha2 = alpha * alpha;
    if (dot(N, L) <= 0.f || dot(N, V) <= 0.f) return 0.f;
    const float nd[2] = {fabsf(dot(N, L)), fabsf(dot(N, V))};
    float res = 1.f;
#pragma unroll
    for (int i = 0; i < 2; ++i) res = res * (2.f * nd[i] / (nd[i] + sqrtf(alpha2 + (1.f - alpha2) * nd[i] * nd[i])));
    return res;
}
RTB_HD float fresnel(float f0, float p5) { return f0 + (1.f - f0) * p5; }   // vector.h:460-471, p5 = (float)pow(1-VdotH,5)

struct Bounce {   // what Scene::intersect needs after its child returns (scene.cpp:154)
    V3 emission;  // intersection.color before the +=
    float coeff;  // 1 / pdf
    V3 material;
    float cosine; // dot(dir, shading_normal)
    float alpha;  // obj.mesh->material.alpha
};

struct PathCounters {
    unsigned long long rays, light_queries;
    Counters tests;   // closest-hit traversal (scene BVH)
    Counters light;   // all-hits light queries (light BVH)
};

// Scene::intersect (scene.cpp:71-156) for one camera sample: returns intersection.color of the root call.
template <int CULL, bool COUNT>
RTB_HD V3 trace_path(const DScene &sc, Ray r, Rng &g, PathCounters *pc) {
    Bounce st[kMaxRayDepth];
    int n = 0;   // vertices that returned a successful Intersection so far
    const int depth = sc.ray_depth;
    for (int k = 0; k < depth; ++k) {
        Hit h;
        pc->rays += 1;
        const bool ok = closest_hit<CULL, COUNT>(sc.bvh, r, h, COUNT ? &pc->tests : nullptr);
        if (!(ok && h.t < sc.max_distance)) break;   // miss: `intersection.object_id >= objects.size()` -> unsuccessful
        const float4 q2 = ldg(sc.bvh.tris + (size_t) h.id * kTriStride + 2);
        V3 Ng = V3{q2.y, q2.z, q2.w};
        const V3 rd = V3{r.d[0], r.d[1], r.d[2]};
        const bool inside = dot(rd, Ng) > 0.f;   // primitive.cpp:51-54
        if (inside) Ng = -Ng;
        const Shade s = load_shade(sc, h.id);
        const Material m = load_material(sc, s.mesh);
        Bounce &b = st[n];
        ++n;
        b.emission = textured_factor(sc, s, m.emission_i, -1, m.emission, h.u, h.v);
        b.coeff = 0.f;
        // the last vertex: its child is `power <= 0 -> return {}`; only the direction sample touches the stream
        const V3 pos = V3{r.o[0], r.o[1], r.o[2]} + rd * h.t;
        V3 Ns = shading_normal(sc, h.id, s, m, h.u, h.v);
        if (inside) Ns = -Ns;
        float rough2, metallic;
        metallic_roughness(sc, s, m, h.u, h.v, rough2, metallic);
        rough2 = std_max(kRoughness2Limit, rough2);
        const V3 eye = -rd;
        const V3 dir = sample_direction(sc, pos, Ns, eye, rough2, g);
        if (k + 1 >= depth) break;
        if (dot(dir, Ns) <= 0.f) {
            if (dot(dir, Ng) <= 0.f) break;
            Ns = Ng;
        }
        const float pdf = pdf_direction<COUNT>(sc, pos, Ns, eye, rough2, dir, COUNT ? &pc->light : nullptr, &pc->light_queries);
        if (pdf <= 0.f || pdf != pdf) break;
        // the factors of scene.cpp:134-154, evaluated now (they only depend on this vertex)
        const V3 half = normal(dir - rd);
        const float vis = smith_g(rough2, Ns, eye, dir) * (1.f / (4.f * fabsf(dot(Ns, rd)) * fabsf(dot(Ns, dir))));
        const float spec = ggx_d(rough2, Ns, half) * vis;
        const float VdotH = fabsf(dot(eye, half));
        const V3 base = textured_factor(sc, s, m.base_i, m.fused, m.base, h.u, h.v);
        const float p5 = pow5_via_double(1.f - VdotH);
        const V3 metal = spec * V3{fresnel(base.x, p5), fresnel(base.y, p5), fresnel(base.z, p5)};
        const V3 diffuse = base * kInvPi;
        const float dsc = fresnel(0.04f, p5);
        const V3 dielectric = diffuse * (1.f - dsc) + (V3{1.f, 1.f, 1.f} * spec) * dsc;
        b.material = dielectric * (1.f - metallic) + metal * metallic;
        b.coeff = 1.f / pdf;
        b.cosine = dot(dir, Ns);
        b.alpha = m.alpha;
        // Ray reflect_ray(pos + dir * step, dir)
        const V3 o2 = pos + dir * kStep;
        make_ray(r, o2.x, o2.y, o2.z, dir.x, dir.y, dir.z);
    }
    if (n == 0) return V3{sc.bg_color[0], sc.bg_color[1], sc.bg_color[2]};
    // unwind: vertex n-1 had no successful child; every earlier one adds child * coeff * material * cos * alpha
    V3 L = st[n - 1].emission;
    for (int j = n - 2; j >= 0; --j) {
        const Bounce &b = st[j];
        L = b.emission + (((L * b.coeff) * b.material) * b.cosine) * b.alpha;
    }
    return L;
}

// Camera::cast_in_pixel (camera.cpp:49-62) + Ray::Ray
// camera_dir: the un-normalised direction handed to Ray::Ray
RTB_HD V3 camera_dir(const DScene &sc, int px, int py, int width, int height, float ox, float oy) {
    const float tx = (2.f * ((float) px + 0.5f + ox) / (float) width - 1.f) * sc.cam_tan_half_fov[0];
    const float ty = -(2.f * ((float) py + 0.5f + oy) / (float) height - 1.f) * sc.cam_tan_half_fov[1];
    const float tz = 1.f;
    V3 d = V3{0.f, 0.f, 0.f};
    d = d + tx * V3{sc.cam_axes[0], sc.cam_axes[1], sc.cam_axes[2]};
    d = d + ty * V3{sc.cam_axes[3], sc.cam_axes[4], sc.cam_axes[5]};
    d = d + tz * V3{sc.cam_axes[6], sc.cam_axes[7], sc.cam_axes[8]};
    return d;
}
RTB_HD void camera_ray(const DScene &sc, int px, int py, int width, int height, float ox, float oy, Ray &r) {
    const V3 d = camera_dir(sc, px, py, width, height, ox, oy);
    make_ray(r, sc.cam_pos[0], sc.cam_pos[1], sc.cam_pos[2], d.x, d.y, d.z);
}

// One pixel of Scene::render's first loop (scene.cpp:32-43): all samples of the pixel, one sequential stream.
template <int CULL, bool COUNT>
RTB_HD V3 render_pixel(const DScene &sc, int px, int py, int width, int height, int samples, uint32_t pixel0_seed, PathCounters *pc) {
    Rng g;
    const uint64_t seed = (uint64_t) py * (uint64_t) width + (uint64_t) px;
    rng_seed(g, seed != 0 ? seed : (uint64_t) pixel0_seed);
    V3 sum = V3{0.f, 0.f, 0.f};
    for (int s = 0; s < samples; ++s) {
        const float ox = rng_pixel_offset(g);
        const float oy = rng_pixel_offset(g);
        Ray r;
        camera_ray(sc, px, py, width, height, ox, oy, r);
        const V3 c = trace_path<CULL, COUNT>(sc, r, g, pc);
        sum = sum + c;
    }
    return sum;
}

// The resolve loop of Scene::render (scene.cpp:54-63): x (1/samples), ACES, pow(1/2.2f), 8-bit
RTB_HD float aces_channel(float x) {   // vector.h:400-407 + saturate = clamp(v,0,1) = max(min(v,1),0)
    const float num = x * (2.51f * x + 0.03f);
    const float den = x * (2.43f * x + 0.59f) + 0.14f;
    const float q = num / den;
    return std_max(std_min(q, 1.f), 0.f);
}
RTB_HD uint8_t resolve_channel(float sum, float normalizer) {
    float c = sum * normalizer;
    c = aces_channel(c);
    c = glibc_powf_pos(c, 1.f / 2.2f);
    const float v = std_max(std_min(c * 255.f, 255.f), 0.f);   // normal_to_ch8bit (vector.h:222-224)
    if (v != v) return 0;
    return (uint8_t) roundf(v);
}

}  // namespace rtb
