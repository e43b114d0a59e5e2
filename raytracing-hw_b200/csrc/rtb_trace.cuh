// Persistent-warp closest-hit engine (device only, sm_100a): the ray-queue half of the wavefront integrator and the
// body of the batched rtb_closest_hit kernels.
//
// Why (ncu, profiles/r01*_lines.txt): with one ray per thread the warps ran at ~5 active threads per instruction.
// Two causes: (1) the rays of a warp need very different numbers of steps, so the warp lives as long as its
// longest ray; (2) lanes in the node phase and lanes in the triangle phase serialise.  Here
//   * every lane owns one ray at a time and REFILLS itself from a global queue (warp-aggregated atomicAdd) as soon
//     as its ray is finished -- no tail;
//   * each warp iteration runs ONE phase chosen by vote over the lanes' wishes (refill / node step / triangle step);
//     the phase body is predicated, control flow is warp-uniform (__ballot_sync), so lanes wanting the same phase
//     always execute it together.
// The per-ray arithmetic is that of rtb_traverse.cuh (same functions), so results are bit-identical to
// closest_hit_global: same visit order per ray, global-best culling.
#pragma once
#include "rtb_traverse.cuh"

namespace rtb {

struct TracePolicy;
TracePolicy trace_policy();   // rtb_kernels.cu

struct QueueCounters {
    unsigned long long box, tri, rays;
};

// Phase vote of a warp iteration: refill when at least `refill_min` lanes are idle (or nothing else is runnable);
// otherwise the node phase when lanes_in_node_phase * w_node >= lanes_in_triangle_phase * w_leaf, else the triangle
// phase.  A node step costs ~2.5x the instructions of a triangle step, so waiting triangle lanes are cheaper to serve
// early: the default weights come from the sweep in profiles/ (RTB_TRACE_POLICY="refill,w_node,w_leaf" overrides).
struct TracePolicy {
    int refill_min, w_node, w_leaf;
};

// Fetch(idx, Ray&, tag&) loads queue entry `idx` (normalising as Ray::Ray does) and returns false when the entry holds
// no ray (a finished slot of the wavefront integrator); Emit(tag, ray, hit) stores the result of the entry's ray.
template <bool COUNT, class Fetch, class Emit>
__device__ __forceinline__ void trace_persistent(const DBvh &b, unsigned n, unsigned *queue_head, Fetch fetch, Emit emit, QueueCounters *cnt, const TracePolicy pol) {
    const unsigned FULL = 0xffffffffu;
    const int lane = threadIdx.x & 31;
    int s_link[kMaxStack + 1];
    float s_m[kMaxStack + 1];
    s_link[0] = kDoneLink;
    s_m[0] = -1.f;
    int sp = 1;
    int link = kDoneLink;          // >= 0 internal node pending; leaf link: triangles [tri_i, tri_end) pending; kDoneLink: no ray
    int tri_i = 0, tri_end = 0;
    unsigned tag = 0;
    bool have_ray = false, exhausted = false;
    float omax = 0.f;
    Ray r;
    Hit best;
    float cull = 0.f;
    best.t = kFar; best.u = 0.f; best.v = 0.f; best.id = -1;
    for (int k = 0; k < 3; ++k) { r.o[k] = 0.f; r.d[k] = 0.f; r.inv[k] = 0.f; }

    auto set_link = [&](int l) {
        link = l;
        if (l < 0 && l != kDoneLink) {
            int first, count;
            decode_leaf(b, l, first, count);
            tri_i = first;
            tri_end = first + count;
        }
    };
    auto pop = [&]() {
        float m;
        int l;
        do {
            --sp;
            l = s_link[sp];
            m = s_m[sp];
        } while (m > cull);      // the sentinel carries m = -1
        if (l == kDoneLink) sp = 1;
        set_link(l);
    };

    while (true) {
        const bool wantR = (link == kDoneLink) && !exhausted;
        const bool wantN = link >= 0;
        const bool wantL = (link < 0) && (link != kDoneLink);
        const unsigned mR = __ballot_sync(FULL, wantR), mN = __ballot_sync(FULL, wantN), mL = __ballot_sync(FULL, wantL);
        if ((mR | mN | mL) == 0u) break;
        const int cR = __popc(mR), cN = __popc(mN), cL = __popc(mL);
        if (cR >= pol.refill_min || (cN == 0 && cL == 0)) {
            // ---- refill phase: store finished results, claim new rays
            const int leader = __ffs(mR) - 1;
            unsigned base = 0;
            if (lane == leader) base = atomicAdd(queue_head, (unsigned) cR);
            base = __shfl_sync(FULL, base, leader);
            if (wantR) {
                if (have_ray) emit(tag, r, best);
                const unsigned idx = base + (unsigned) __popc(mR & ((1u << lane) - 1u));
                if (idx < n) {
                    have_ray = fetch(idx, r, tag);
                    if (have_ray) {
                        best.t = kFar; best.u = 0.f; best.v = 0.f; best.id = -1;
                        omax = ray_omax(r);
                        cull = cull_bound(kFar, omax);
                        float m0;
                        if (COUNT) cnt->box += 1;
                        cnt->rays += 1;
                        if (box_test_t(r, b.root_box[0], b.root_box[1], b.root_box[2], b.root_box[3], b.root_box[4], b.root_box[5], m0))
                            set_link(b.root_link);
                    }
                } else {
                    have_ray = false;
                    exhausted = true;
                }
            }
        } else if (cN * pol.w_node >= cL * pol.w_leaf) {
            // ---- node phase
            if (wantN) {
                const DNode *np = b.nodes + link;
                const float4 na = __ldg(&np->a), nb = __ldg(&np->b), nc = __ldg(&np->c);
                const int4 nd = __ldg(&np->d);
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
                    set_link(lA);
                } else if (vA || vB) {
                    set_link(vA ? lA : lB);
                } else {
                    pop();
                }
            }
        } else {
            // ---- triangle phase: one triangle per lane
            if (wantL) {
                const float4 *tp = b.tris + (size_t) tri_i * 3;
                const float4 q0 = __ldg(tp), q1 = __ldg(tp + 1), q2 = __ldg(tp + 2);
                float t, u, v;
                if (COUNT) cnt->tri += 1;
                if (tri_test_flat(r, q0, q1, q2, t, u, v) && t < best.t) {
                    best.t = t;
                    best.u = u;
                    best.v = v;
                    best.id = tri_i;
                    cull = cull_bound(t, omax);
                }
                ++tri_i;
                if (tri_i == tri_end) pop();
            }
        }
    }
}

}  // namespace rtb
