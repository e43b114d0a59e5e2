/* TEST INFRASTRUCTURE ONLY -- see twin.h.  Plain-C restatement of the reference's traversal hot path, written as
 * the reference is written (recursive, one ray at a time), NOT as the CUDA kernels are (iterative, packed nodes).
 * Compile with -ffp-contract=off and no -ffast-math: the reference's Release build targets the x86-64 baseline
 * (SSE2), where no fused multiply-add exists, so every float expression here rounds after each operation.
 *
 * Parity: pinned against the reference itself (oracle/_ref) and tests/golden/ -- see twin.h.
 */
#include "twin.h"

#include <math.h>
#include <omp.h>
#include <string.h>

/* ------------------------------------------------------------------------------------------------ vectors
 * dot: s = 0; s += a[i]*b[i]            (src/utils/vector.h:365-370)
 * cross: a1*b2 - b1*a2, ...             (src/utils/vector.h:372-378)
 * a - b == a + (-b)                     (src/utils/vector.h:326-328)                                              */
static float dot3(const float *a, const float *b) {
    float s = 0.f;
    for (int i = 0; i < 3; ++i) s += a[i] * b[i];
    return s;
}
static void cross3(const float *a, const float *b, float *o) {
    o[0] = a[1] * b[2] - b[1] * a[2];
    o[1] = a[2] * b[0] - b[2] * a[0];
    o[2] = a[0] * b[1] - b[0] * a[1];
}
static void sub3(const float *a, const float *b, float *o) {
    for (int i = 0; i < 3; ++i) o[i] = a[i] + (-b[i]);
}
static float length3(const float *a) { return sqrtf(dot3(a, a)); }

typedef struct {
    float origin[3], direction[3], inv_direction[3];
} ray_t;

/* Ray::Ray (src/geometry/primitive.cpp:12-15); normalize multiplies by 1.f/length (src/utils/vector.h:191-197) */
static void ray_init(ray_t *r, const float *p, const float *d) {
    const float mod = 1.f / length3(d);
    for (int i = 0; i < 3; ++i) {
        r->origin[i] = p[i];
        r->direction[i] = d[i] * mod;
    }
    for (int i = 0; i < 3; ++i) r->inv_direction[i] = 1.f / r->direction[i];
}

/* AABB::intersect (src/geometry/primitive.cpp:146-208).  Returns 1 and the entry distance, or 0. */
static int aabb_intersect(const float *box6, const ray_t *r, float *dist, uint64_t *nbox) {
    const float *mn = box6, *mx = box6 + 3;
    enum { Left = 0, Mid = 1, Right = 2 };
    int inside = 1, quadrant[3];
    float plane[3] = {0.f, 0.f, 0.f};
    if (nbox) ++*nbox;
    for (int i = 0; i < 3; ++i) {
        if (r->origin[i] < mn[i]) {
            quadrant[i] = Left;
            plane[i] = mn[i];
            inside = 0;
        } else if (r->origin[i] > mx[i]) {
            quadrant[i] = Right;
            plane[i] = mx[i];
            inside = 0;
        } else {
            quadrant[i] = Mid;
        }
    }
    if (inside) {
        *dist = 0.f;
        return 1;
    }
    float maxT[3];
    for (int i = 0; i < 3; ++i) {
        if (quadrant[i] != Mid && r->direction[i] != 0.f)
            maxT[i] = (plane[i] - r->origin[i]) * r->inv_direction[i];
        else
            maxT[i] = -1.f;
    }
    int which = 0;
    for (int i = 1; i < 3; ++i)
        if (maxT[which] < maxT[i]) which = i;
    if (maxT[which] < 0.f) return 0;
    float coord[3];
    for (int i = 0; i < 3; ++i) {
        if (which != i) {
            coord[i] = r->origin[i] + maxT[which] * r->direction[i];
            if (coord[i] < mn[i] || coord[i] > mx[i]) return 0;
        } else {
            coord[i] = plane[i];
        }
    }
    float e[3];
    sub3(coord, r->origin, e);
    *dist = length3(e);
    return 1;
}

typedef struct {
    int successful;
    float distance, u, v;
    int64_t object_id;
} hit_t;

/* Primitive::intersect (src/geometry/primitive.cpp:17-57); `-1e-6 < det && det < 1e-6` compares in double */
static hit_t tri_intersect(const float *pos9, const ray_t *r, uint64_t *ntri) {
    hit_t h = {0, 0.f, 0.f, 0.f, -1};
    const float *p0 = pos9, *U = pos9 + 3, *V = pos9 + 6;
    float P[3], s[3], Q[3];
    if (ntri) ++*ntri;
    cross3(r->direction, V, P);
    const float det = dot3(U, P);
    if (-1e-6 < (double) det && (double) det < 1e-6) return h;
    const float inv_det = 1.f / det;
    sub3(r->origin, p0, s);
    const float u = inv_det * dot3(s, P);
    if (u < 0 || u > 1) return h;
    cross3(s, U, Q);
    const float v = inv_det * dot3(r->direction, Q);
    if (v < 0 || u + v > 1) return h;
    const float t = inv_det * dot3(V, Q);
    if (t < 0.f) return h;
    h.successful = 1;
    h.distance = t;
    h.u = u;
    h.v = v;
    return h;
}

#define LEFT(b, i) ((b)->links[(size_t) (i) * 5 + 0])
#define RIGHT(b, i) ((b)->links[(size_t) (i) * 5 + 1])
#define SPLIT(b, i) ((b)->links[(size_t) (i) * 5 + 2])
#define FIRST(b, i) ((b)->links[(size_t) (i) * 5 + 3])
#define COUNT(b, i) ((b)->links[(size_t) (i) * 5 + 4])

/* BVH::intersectHelper (src/core/bvh.cpp:177-237): the near child (by the SIGN of direction[split_dim]) is always
 * entered; the far child is skipped only when its box distance exceeds the best hit of the near child's subtree. */
static hit_t closest_helper(const twin_bvh *b, const ray_t *r, int32_t node, uint64_t *cnt) {
    hit_t best = {0, 1e9f, 0.f, 0.f, -1};
    const int32_t sd = SPLIT(b, node);
    /* a leaf made by the count<=4 early-out has split_dim = invalidKey; the reference still evaluates
     * direction[-1] there, but both children are invalid, so the outcome does not depend on it */
    const int left_first = (sd >= 0 && sd <= 2) ? (r->direction[sd] > 0) : 0;
    const int32_t c0 = left_first ? LEFT(b, node) : RIGHT(b, node);
    const int32_t c1 = left_first ? RIGHT(b, node) : LEFT(b, node);
    float d;
    if (c0 >= 0 && aabb_intersect(b->aabb + (size_t) c0 * 6, r, &d, cnt ? &cnt[0] : 0)) {
        const hit_t a = closest_helper(b, r, c0, cnt);
        if (a.successful && a.distance < best.distance) best = a;
    }
    if (c1 >= 0 && aabb_intersect(b->aabb + (size_t) c1 * 6, r, &d, cnt ? &cnt[0] : 0)) {
        if (d > best.distance) {
        } else {
            const hit_t a = closest_helper(b, r, c1, cnt);
            if (a.successful && a.distance < best.distance) best = a;
        }
    }
    const int32_t pc = COUNT(b, node);
    if (pc <= 0) return best;
    const int32_t f = FIRST(b, node);
    for (int32_t i = f; i < f + pc; ++i) {
        hit_t a = tri_intersect(b->tri_pos + (size_t) i * 9, r, cnt ? &cnt[1] : 0);
        if (a.successful && a.distance < best.distance) {
            best = a;
            best.object_id = i;
        }
    }
    return best;
}

double twin_closest_hit(const twin_bvh *b, const float *rays6, size_t n, int32_t *id, float *t, float *u, float *v,
                        uint8_t *inside, uint64_t *counters2, int threads) {
    if (threads <= 0) threads = omp_get_max_threads();
    uint64_t nbox = 0, ntri = 0;
    const double t0 = omp_get_wtime();
#pragma omp parallel for schedule(dynamic, 1024) num_threads(threads) reduction(+ : nbox, ntri)
    for (size_t i = 0; i < n; ++i) {
        ray_t r;
        uint64_t c[2] = {0, 0};
        ray_init(&r, rays6 + i * 6, rays6 + i * 6 + 3);
        hit_t h = {0, 0.f, 0.f, 0.f, -1};
        float d;
        /* BVH::intersect (src/core/bvh.cpp:239-243): root box first */
        if (aabb_intersect(b->aabb, &r, &d, counters2 ? &c[0] : 0)) h = closest_helper(b, &r, 0, counters2 ? c : 0);
        /* `if (intersection.object_id >= objects.size())` -> miss (src/core/scene.cpp:85): a helper result that
         * never saw a triangle keeps object_id = invalidKey */
        const int ok = h.successful && h.object_id >= 0;
        id[i] = ok ? (int32_t) h.object_id : -1;
        if (t) t[i] = ok ? h.distance : 0.f;
        if (u) u[i] = ok ? h.u : 0.f;
        if (v) v[i] = ok ? h.v : 0.f;
        if (inside) {
            uint8_t in = 0;
            if (ok) {
                /* get_geometric_normal = normal(cross(U, V)) (src/geometry/primitive.cpp:77-84);
                 * `if (dot(ray.direction, normal) > 0) inside = true` (src/geometry/primitive.cpp:51-54) */
                const float *p = b->tri_pos + (size_t) h.object_id * 9;
                float ng[3];
                cross3(p + 3, p + 6, ng);
                const float mod = 1.f / length3(ng);
                for (int k = 0; k < 3; ++k) ng[k] *= mod;
                in = dot3(r.direction, ng) > 0 ? 1 : 0;
            }
            inside[i] = in;
        }
        nbox += c[0];
        ntri += c[1];
    }
    if (counters2) {
        counters2[0] = nbox;
        counters2[1] = ntri;
    }
    return omp_get_wtime() - t0;
}

/* BVH::intersectAllHelper (src/core/bvh.cpp:245-273) fused with the accumulation of ManyLightsDistribution::pdf
 * (src/utils/random.cpp:179-188) and LightDistribution::pdfWithHint (src/utils/random.cpp:83-91):
 * hits are visited left subtree, right subtree, own triangles -- the order they are pushed into the vector. */
static void all_helper(const twin_bvh *b, const ray_t *r, int32_t node, const float *dir, float *prob) {
    float d;
    const int32_t l = LEFT(b, node), rt = RIGHT(b, node);
    if (l >= 0 && aabb_intersect(b->aabb + (size_t) l * 6, r, &d, 0)) all_helper(b, r, l, dir, prob);
    if (rt >= 0 && aabb_intersect(b->aabb + (size_t) rt * 6, r, &d, 0)) all_helper(b, r, rt, dir, prob);
    const int32_t pc = COUNT(b, node);
    if (pc <= 0) return;
    const int32_t f = FIRST(b, node);
    for (int32_t i = f; i < f + pc; ++i) {
        const float *p = b->tri_pos + (size_t) i * 9;
        const hit_t a = tri_intersect(p, r, 0);
        if (!a.successful) continue;
        /* triangle_area = length(cross(U,V)) * 0.5f, probability = 1.f / triangle_area (random.cpp:86-88) */
        float c[3], ng[3];
        cross3(p + 3, p + 6, c);
        const float len = length3(c);
        const float area = 0.5f * len;
        const float probability = 1.f / area;
        const float mod = 1.f / len;
        for (int k = 0; k < 3; ++k) ng[k] = c[k] * mod;
        /* the Intersection's normal is flipped toward the ray when inside; abs() absorbs the sign */
        *prob += fabsf(probability * (a.distance * a.distance) / dot3(ng, dir));
    }
}

void twin_light_pdf(const twin_bvh *b, const float *rays6, size_t n, float *pdf) {
    for (size_t i = 0; i < n; ++i) {
        ray_t r;
        ray_init(&r, rays6 + i * 6, rays6 + i * 6 + 3);
        float prob = 0.f;
        /* pdfWithHint receives the caller's `direction` (not re-normalised), random.cpp:183-186 */
        all_helper(b, &r, 0, rays6 + i * 6 + 3, &prob);
        pdf[i] = prob / (float) b->n_tris;
    }
}

/* ------------------------------------------------------------------------------------------------ <random>
 * std::minstd_rand = linear_congruential_engine<uint_fast32_t, 48271, 0, 2147483647>; seed(s): s mod m, 0 -> 1.
 * generate_canonical<float, 24>: one engine call, (x - 1) / 2147483646.0f (the divisor rounds to 2^31), results
 * >= 1 become nextafter(1, 0) (bits/random.tcc:3349-3380).  uniform_real: u * (b - a) + a.
 * normal_distribution<float>: Marsaglia polar, returns y*mult and caches x*mult (bits/random.tcc:1812-1845).      */
static uint32_t lcg_seed(uint32_t s) {
    uint32_t x = s % 2147483647u;
    return x == 0 ? 1u : x;
}
static uint32_t lcg_next(uint32_t *x) {
    *x = (uint32_t) (((uint64_t) *x * 48271ull) % 2147483647ull);
    return *x;
}
static float canonical(uint32_t *x) {
    float u = (float) (lcg_next(x) - 1u) / 2147483648.f;
    if (u >= 1.f) u = nextafterf(1.f, 0.f);
    return u;
}
uint32_t twin_minstd_first(uint32_t seed) {
    uint32_t x = lcg_seed(seed);
    return lcg_next(&x);
}
float twin_uniform_first(uint32_t seed, float a, float b) {
    uint32_t x = lcg_seed(seed);
    return canonical(&x) * (b - a) + a;
}
void twin_normal_seq(uint32_t seed, int n, float *out) {
    uint32_t x = lcg_seed(seed);
    int have = 0;
    float saved = 0.f;
    for (int i = 0; i < n; ++i) {
        float ret;
        if (have) {
            have = 0;
            ret = saved;
        } else {
            float a, b, r2;
            do {
                a = 2.f * canonical(&x) - 1.f;
                b = 2.f * canonical(&x) - 1.f;
                r2 = a * a + b * b;
            } while (r2 > 1.f || r2 == 0.f);
            const float mult = sqrtf(-2.f * logf(r2) / r2);
            saved = a * mult;
            have = 1;
            ret = b * mult;
        }
        out[i] = ret * 1.f + 0.f;
    }
}

/* resolve (src/core/scene.cpp:54-63): color * (1/samples) -> aces_tonemap (src/utils/vector.h:400-407) ->
 * powf(., 1/2.2f) (scene.h:36) -> normal_to_ch8bit = round(clamp(255 v, 0, 255)) (src/utils/vector.h:222-233) */
void twin_resolve(const float *accum, size_t n_values, int samples, uint8_t *rgb8) {
    const float normalizer = 1.f / (float) samples;
    const float gamma = 1.f / 2.2f;
    for (size_t i = 0; i < n_values; ++i) {
        float x = accum[i] * normalizer;
        const float a = 2.51f, b = 0.03f, c = 2.43f, d = 0.59f, e = 0.14f;
        x = (x * (a * x + b)) / (x * (c * x + d) + e);
        x = fmaxf(fminf(x, 1.f), 0.f);
        x = powf(x, gamma);
        const float v = fmaxf(fminf(x * 255.f, 255.f), 0.f);
        rgb8[i] = (uint8_t) roundf(v);
    }
}
