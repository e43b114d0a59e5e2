// SIMT-convergent closest-hit traversal (RTB_CULL_GLOBAL), sm_100a.
//
// ncu evidence that shaped it (profiles/r01_*.txt): the first kernel (rtb_device.cuh: closest_hit_ordered, kept for
// RTB_CULL_REFERENCE) ran at 5.2 active threads per warp instruction -- box tests at ~8, triangle tests at ~2, the pop
// loop at 1.5 -- and was issue-bound (71 % issue-active).  Here the warp runs three phases in lock step:
//   node phase : every lane that holds an internal node fetches it (4 x LDG.128), runs both ray/box tests
//                branch-free, and descends / pushes / pops; lanes that reached a leaf wait at the loop exit;
//   leaf phase : every lane that holds a leaf tests one triangle per iteration (3 x LDG.128, straight-line
//                Moller-Trumbore with a single IEEE division);
//   pop        : next deferred sibling whose entry distance does not exceed the best hit.
// Arithmetic contract unchanged (see rtb_device.cuh): the hit / miss decisions are restatements of AABB::intersect
// (src/geometry/primitive.cpp:146-208) and Primitive::intersect (src/geometry/primitive.cpp:17-57) with identical
// rounding; only control flow is predicated.
//
// Culling (RTB_CULL_GLOBAL only).  The reference culls a box when its entry distance
// length(hit point - origin) (primitive.cpp:207) exceeds the best hit.  Global-best culling may cull LESS than that
// without changing the answer (the triangles tested stay a superset, in the same order), so the entry distance is
// not evaluated at all: a box is culled when the ray parameter of its entry plane, tmax, exceeds
//     cull_bound(best.t) = best.t * (1 + 1e-5) + 1e-6 * max|origin_i|,
// which implies the reference's own distance exceeds best.t: the reference's distance differs from tmax by the
// rounding of |d| (~1e-7 relative) and by the cancellation in (o + tmax*d) - o (<= 2 ulp of max(|o_i|, |c_i|) per
// axis).  box_test_sq / sq_threshold (the exact squared-distance form) remain for reference and tests.
#pragma once
#include "rtb_device.cuh"

namespace rtb {

constexpr int kDoneLink = (int) 0x8FFFFFFFu;   // stack-bottom sentinel: "big leaf" index 2^28-1, never allocated

// largest float m >= 0 with sqrtf_rn(m) <= b, for b >= 0 (sqrtf is monotonic and correctly rounded)
RTB_HD float sq_threshold(float b) {
    float c = b * b;
    while (sqrtf(c) > b) c = u2f(f2u(c) - 1u);
    while (true) {
        const float n = u2f(f2u(c) + 1u);
        if (!(sqrtf(n) <= b)) break;
        c = n;
    }
    return c;
}

// AABB::intersect, predicated.  hit <=> the reference returns a value; m = the SQUARED entry distance
// (((0 + ex*ex) + ey*ey) + ez*ez of primitive.cpp:207 before the sqrt; 0 when the origin is inside).
RTB_HD bool box_test_sq(const Ray &r, float mnx, float mny, float mnz, float mxx, float mxy, float mxz, float &m) {
    const bool lo0 = r.o[0] < mnx, hi0 = r.o[0] > mxx;
    const bool lo1 = r.o[1] < mny, hi1 = r.o[1] > mxy;
    const bool lo2 = r.o[2] < mnz, hi2 = r.o[2] > mxz;
    const float pl0 = lo0 ? mnx : mxx, pl1 = lo1 ? mny : mxy, pl2 = lo2 ? mnz : mxz;
    const bool out0 = lo0 || hi0, out1 = lo1 || hi1, out2 = lo2 || hi2;   // quadrant != Mid
    const bool inside = !(out0 || out1 || out2);
    const float t0 = (out0 && r.d[0] != 0.f) ? (pl0 - r.o[0]) * r.inv[0] : -1.f;
    const float t1 = (out1 && r.d[1] != 0.f) ? (pl1 - r.o[1]) * r.inv[1] : -1.f;
    const float t2 = (out2 && r.d[2] != 0.f) ? (pl2 - r.o[2]) * r.inv[2] : -1.f;
    int which = 0;
    float tmax = t0;
    if (tmax < t1) { which = 1; tmax = t1; }
    if (tmax < t2) { which = 2; tmax = t2; }
    const float c0 = r.o[0] + tmax * r.d[0];
    const float c1 = r.o[1] + tmax * r.d[1];
    const float c2 = r.o[2] + tmax * r.d[2];
    const bool bad0 = (c0 < mnx) || (c0 > mxx);
    const bool bad1 = (c1 < mny) || (c1 > mxy);
    const bool bad2 = (c2 < mnz) || (c2 > mxz);
    const bool bad = (which != 0 && bad0) || (which != 1 && bad1) || (which != 2 && bad2) || (tmax < 0.f);
    const float ex = (which == 0 ? pl0 : c0) - r.o[0];
    const float ey = (which == 1 ? pl1 : c1) - r.o[1];
    const float ez = (which == 2 ? pl2 : c2) - r.o[2];
    float s = ex * ex;
    s = s + ey * ey;
    s = s + ez * ez;
    m = inside ? 0.f : s;
    return inside || !bad;
}

// AABB::intersect, predicated, without the entry distance.  hit <=> the reference returns a value; tmax = the ray
// parameter of the candidate plane (primitive.cpp:176-190), -1 when the origin is inside the box.
// NZ: the caller guarantees that no component of the direction is zero (the `dir[i] != 0.` guards of primitive.cpp:178
// are then always true and are not evaluated).
template <bool NZ = false>
RTB_HD bool box_test_t(const Ray &r, float mnx, float mny, float mnz, float mxx, float mxy, float mxz, float &tmax) {
    const bool lo0 = r.o[0] < mnx, hi0 = r.o[0] > mxx;
    const bool lo1 = r.o[1] < mny, hi1 = r.o[1] > mxy;
    const bool lo2 = r.o[2] < mnz, hi2 = r.o[2] > mxz;
    const float pl0 = lo0 ? mnx : mxx, pl1 = lo1 ? mny : mxy, pl2 = lo2 ? mnz : mxz;
    const bool out0 = lo0 || hi0, out1 = lo1 || hi1, out2 = lo2 || hi2;   // quadrant != Mid
    const float t0 = (out0 && (NZ || r.d[0] != 0.f)) ? (pl0 - r.o[0]) * r.inv[0] : -1.f;
    const float t1 = (out1 && (NZ || r.d[1] != 0.f)) ? (pl1 - r.o[1]) * r.inv[1] : -1.f;
    const float t2 = (out2 && (NZ || r.d[2] != 0.f)) ? (pl2 - r.o[2]) * r.inv[2] : -1.f;
    const bool w1 = t0 < t1;                 // whichPlane = 1 (unless 2 wins)
    float tm = w1 ? t1 : t0;
    const bool w2 = tm < t2;                 // whichPlane = 2
    tm = w2 ? t2 : tm;
    const float c0 = r.o[0] + tm * r.d[0];
    const float c1 = r.o[1] + tm * r.d[1];
    const float c2 = r.o[2] + tm * r.d[2];
    const bool bad0 = (c0 < mnx) || (c0 > mxx);
    const bool bad1 = (c1 < mny) || (c1 > mxy);
    const bool bad2 = (c2 < mnz) || (c2 > mxz);
    // the coordinate of whichPlane itself is not checked (primitive.cpp:195-203)
    const bool bad = ((w1 || w2) && bad0) || ((!w1 || w2) && bad1) || (!w2 && bad2) || (tm < 0.f);
    const bool inside = !(out0 || out1 || out2);
    tmax = inside ? -1.f : tm;
    return inside || !bad;
}
// AABB::intersect once more, in the form the queue kernels run (fewest compare / select instructions; the first form
// kept ~20 predicates alive per node and ptxas parked them in a bit mask, profiles/r01h_k_wf_trace_lines.txt).  Same decisions as
// box_test_t for EVERY input, by these three observations:
//  (1) Mid axes carry the sentinel -0.f instead of the reference's -1.  A sentinel only matters when no outside axis
//      has a positive parameter (an outside axis' parameter is never +-0: |pl - o| >= the smallest subnormal and
//      |inv| >= 1 - 2^-23 for a normalised direction).  Then the reference rejects unless the origin is inside
//      (tmax < 0, primitive.cpp:193); here tm = -0, so every c_j = o_j + (-0 * d_j) = o_j exactly, which is in
//      bounds exactly when axis j is Mid: all Mid (origin inside, primitive.cpp:172) -> accepted; an outside axis j
//      is never whichPlane (a negative parameter loses against -0) -> rejected by its own bounds check.  The `inside`
//      flag is therefore not computed at all.  tmax is -0 for an origin inside the box: never culled, like -1.
//  (2) The `dir[i] != 0.` guards of primitive.cpp:178 are dropped: for a zero component of an outside axis the
//      product is +-inf.  +inf becomes tm and sends every other coordinate with a non-zero direction to +-inf (out of
//      bounds; there is one, the direction is normalised): rejected.  -inf never wins, and c_i = o_i + tm * 0 = o_i
//      is outside: rejected.  The reference rejects both (t_i = -1, then either tmax < 0 or c_i = o_i fails).
//  (3) An outside axis with a negative parameter moves away from its slab: c_i <= o_i < mn_i (or >= o_i > mx_i) by
//      monotonic rounding, so it is rejected here as in the reference whichever plane is whichPlane.
// tests/test_devsim.py::test_lane_form3_* pin this against the reference on random, axis-parallel and grid-aligned rays.
RTB_HD bool box_test_z(const Ray &r, float mnx, float mny, float mnz, float mxx, float mxy, float mxz, float &tmax) {
    const bool lo0 = r.o[0] < mnx, lo1 = r.o[1] < mny, lo2 = r.o[2] < mnz;
    const bool out0 = lo0 || (r.o[0] > mxx), out1 = lo1 || (r.o[1] > mxy), out2 = lo2 || (r.o[2] > mxz);
    const float t0 = out0 ? ((lo0 ? mnx : mxx) - r.o[0]) * r.inv[0] : -0.f;
    const float t1 = out1 ? ((lo1 ? mny : mxy) - r.o[1]) * r.inv[1] : -0.f;
    const float t2 = out2 ? ((lo2 ? mnz : mxz) - r.o[2]) * r.inv[2] : -0.f;
    const bool w1 = t0 < t1;                 // whichPlane = 1 (unless 2 wins)
    float tm = w1 ? t1 : t0;
    const bool w2 = tm < t2;                 // whichPlane = 2
    tm = w2 ? t2 : tm;
    const float c0 = r.o[0] + tm * r.d[0];
    const float c1 = r.o[1] + tm * r.d[1];
    const float c2 = r.o[2] + tm * r.d[2];
    const bool bad0 = (c0 < mnx) || (c0 > mxx);
    const bool bad1 = (c1 < mny) || (c1 > mxy);
    const bool bad2 = (c2 < mnz) || (c2 > mxz);
    tmax = tm;
    return !(((w1 || w2) && bad0) || ((!w1 || w2) && bad1) || (!w2 && bad2) || (tm < 0.f));
}
// see the header comment: tmax > cull_bound(best.t, omax) implies the reference's entry distance > best.t
RTB_HD float ray_omax(const Ray &r) { return fmaxf(fmaxf(fabsf(r.o[0]), fabsf(r.o[1])), fabsf(r.o[2])); }
RTB_HD float cull_bound(float best_t, float omax) { return best_t * 1.00001f + 1e-6f * omax; }

// Primitive::intersect, straight-line: one division, rejections folded into a flag with the reference's NaN
// behaviour (every rejection is a comparison that is false for NaN).
RTB_HD bool tri_test_flat(const Ray &r, const float4 q0, const float4 q1, const float4 q2, float &t, float &u, float &v) {
    const float Ux = q0.w, Uy = q1.x, Uz = q1.y;
    const float Vx = q1.z, Vy = q1.w, Vz = q2.x;
    const float Px = r.d[1] * Vz - Vy * r.d[2];
    const float Py = r.d[2] * Vx - Vz * r.d[0];
    const float Pz = r.d[0] * Vy - Vx * r.d[1];
    float det = Ux * Px;
    det = det + Uy * Py;
    det = det + Uz * Pz;
    const bool parallel = fabsf(det) <= u2f(0x358637bdu);
    const float inv_det = 1.f / det;
    const float sx = r.o[0] - q0.x, sy = r.o[1] - q0.y, sz = r.o[2] - q0.z;
    float su = sx * Px;
    su = su + sy * Py;
    su = su + sz * Pz;
    u = inv_det * su;
    const float Qx = sy * Uz - Uy * sz;
    const float Qy = sz * Ux - Uz * sx;
    const float Qz = sx * Uy - Ux * sy;
    float sv = r.d[0] * Qx;
    sv = sv + r.d[1] * Qy;
    sv = sv + r.d[2] * Qz;
    v = inv_det * sv;
    float st = Vx * Qx;
    st = st + Vy * Qy;
    st = st + Vz * Qz;
    t = inv_det * st;
    const bool reject = parallel || (u < 0.f) || (u > 1.f) || (v < 0.f) || (u + v > 1.f) || (t < 0.f);
    return !reject;
}

// BVH::intersect (src/core/bvh.cpp:239-243) with global-best culling: same visit order as the reference (near child
// by the sign of direction[split_dim], deferred far sibling), every box culled against the best hit so far.
template <bool COUNT>
RTB_HD bool closest_hit_global(const DBvh &b, const Ray &r, Hit &best, Counters *cnt) {
    best.t = kFar;
    best.id = -1;
    best.u = 0.f;
    best.v = 0.f;
    float m0;
    if (COUNT) cnt->box += 1;
    if (!box_test_t(r, b.root_box[0], b.root_box[1], b.root_box[2], b.root_box[3], b.root_box[4], b.root_box[5], m0)) return false;

    int s_link[kMaxStack + 1];
    float s_m[kMaxStack + 1];
    s_link[0] = kDoneLink;
    s_m[0] = -1.f;
    int sp = 1;
    const float omax = ray_omax(r);
    float cull = cull_bound(kFar, omax);   // image of best.t for the entry-plane parameter
    int link = b.root_link;

    while (link != kDoneLink) {
        // ---- node phase
        while (link >= 0) {
            const DNode *np = b.nodes + link;
            const float4 na = ldg(&np->a), nb = ldg(&np->b), nc = ldg(&np->c);
            const int4 nd = ldg(&np->d);
            float ml, mr;
            const bool hl = box_test_t(r, na.x, na.y, na.z, na.w, nb.x, nb.y, ml);
            const bool hr = box_test_t(r, nb.z, nb.w, nc.x, nc.y, nc.z, nc.w, mr);
            if (COUNT) cnt->box += 2;
            const float dax = nd.z == 0 ? r.d[0] : (nd.z == 1 ? r.d[1] : r.d[2]);
            const bool left_first = dax > 0.f;   // bvh.cpp:183
            const int lA = left_first ? nd.x : nd.y, lB = left_first ? nd.y : nd.x;
            const float mA = left_first ? ml : mr, mB = left_first ? mr : ml;
            const bool vA = (left_first ? hl : hr) && !(mA > cull);
            const bool vB = (left_first ? hr : hl) && !(mB > cull);
            if (vA && vB) {
                s_link[sp] = lB;
                s_m[sp] = mB;
                ++sp;
                link = lA;
            } else if (vA || vB) {
                link = vA ? lA : lB;
            } else {
                float m;
                do {
                    --sp;
                    link = s_link[sp];
                    m = s_m[sp];
                } while (m > cull);   // the sentinel carries m = -1: never culled
            }
        }
        if (link == kDoneLink) break;
        // ---- leaf phase
        {
            int first, count;
            decode_leaf(b, link, first, count);
            for (int i = first; i < first + count; ++i) {
                const float4 *tp = b.tris + (size_t) i * kTriStride;
                const float4 q0 = ldg(tp), q1 = ldg(tp + 1), q2 = ldg(tp + 2);
                float t, u, v;
                if (COUNT) cnt->tri += 1;
                if (tri_test_flat(r, q0, q1, q2, t, u, v) && t < best.t) {
                    best.t = t;
                    best.u = u;
                    best.v = v;
                    best.id = i;
                    cull = cull_bound(t, omax);
                }
            }
        }
        // ---- pop
        {
            float m;
            do {
                --sp;
                link = s_link[sp];
                m = s_m[sp];
            } while (m > cull);
        }
    }
    return best.id >= 0;
}

template <int CULL, bool COUNT>
RTB_HD bool closest_hit(const DBvh &b, const Ray &r, Hit &best, Counters *cnt) {
    if (CULL == 0) return closest_hit_ordered<0, COUNT>(b, r, best, cnt);
    return closest_hit_global<COUNT>(b, r, best, cnt);
}

}  // namespace rtb
