// Persistent-warp closest-hit engine (device only, sm_100a): the ray-queue half of the wavefront integrator and the
// body of the batched rtb_closest_hit kernels.
//
// Why (ncu, profiles/r01*_lines.txt): with one ray per thread the warps ran at ~5 active threads per instruction.
// Two causes: (1) the rays of a warp need very different numbers of steps, so the warp lives as long as its
// longest ray; (2) lanes in the node phase and lanes in the triangle phase serialise.  Here
//   * every lane owns one ray at a time and REFILLS itself from a global queue (warp-aggregated atomicAdd) as soon
//     as its ray is finished;
//   * a STEP serves every lane: pops and leaf decodes at the top (one site each, where the warp is converged), one
//     fetch stage in which a lane loads its 64-byte record -- its node or its triangle -- then the node code for
//     the lanes holding a node and the triangle code for the lanes holding a triangle, both predicated;
//   * the tail of a launch is bounded: rays still running long after the queue is empty are suspended and resumed
//     by the next launch (see trace_persistent).
// The per-ray arithmetic is that of rtb_traverse.cuh, so results are bit-identical to closest_hit_global: same visit
// order per ray, global-best culling.
#pragma once
#include "rtb_traverse.cuh"

// entries of the form-3 stack kept in shared memory per thread (a power of two; 0: none), per kernel: measured in
// profiles/r01m_stack_ring_experiment.txt -- the ring wins on incoherent batches (soup +6 %) and loses on the frame (-3.5 %)
#ifndef RTB_RING_BATCH
#define RTB_RING_BATCH 8
#endif
#ifndef RTB_RING_WAVEFRONT
#define RTB_RING_WAVEFRONT 0
#endif

// Latency experiments of round 2 (each a build-time switch so that variants can be A/B-ed on the box, tools/gpu_run.sh variants):
//   RTB_PREFETCH_LEAF    : while triangle i of a leaf is tested, prefetch.global.L1 the record of triangle i + 1 (the
//                          triangles of a leaf are contiguous and are all tested: its address is known a step ahead)
//   RTB_PREFETCH_SIBLING : at a push, prefetch the deferred sibling's record
//   RTB_TRACE_STATE_CS   : the queue kernel's path-state traffic (ray in, hit out) with evict-first (.cs) hints, so that
//                          streaming state stops displacing BVH records from the L1
#ifndef RTB_PREFETCH_LEAF
#define RTB_PREFETCH_LEAF 0
#endif
#ifndef RTB_PREFETCH_SIBLING
#define RTB_PREFETCH_SIBLING 0
#endif
#ifndef RTB_TRACE_STATE_CS
#define RTB_TRACE_STATE_CS 0
#endif
//   RTB_FETCH_HINTS      : L1 policy of the record fetch by kind -- 1: triangle records do not allocate in the L1 (a scene's triangles
//                          are far more than an L1 holds and are rarely fetched twice by the same SM; the nodes near the root are
//                          fetched by every ray), 2: the same and node records with evict_last
#ifndef RTB_FETCH_HINTS
#define RTB_FETCH_HINTS 0
#endif

namespace rtb {

constexpr int kRingThreads = 128;   // threads per CTA of the kernels that run the form-3 loop

struct TracePolicy;
TracePolicy trace_policy();   // rtb_kernels.cu

struct QueueCounters {
    unsigned long long box, tri;
    unsigned rays;          // per lane and launch
};

// A warp refills from the queue when at least `refill_min` of its lanes are idle (or none is busy); the default comes
// from the sweeps in profiles/ (RTB_TRACE_POLICY="refill,1,1" overrides; the two weights belonged to the first form's
// phase vote and are no longer read).
struct TracePolicy {
    int refill_min, w_node, w_leaf;
};

constexpr int kPopLink = (int) 0x8FFFFFFEu;    // "a pop is pending": big-leaf index 2^28-2, never allocated

// One lane's traversal state and its steps.  RTB_HD so that tests/devsim can run the same state machine on the host.
// States (by `link`): >= 0 internal node to visit; a leaf link while its triangles are being tested (first form: the
// range [tri_i, tri_end); form 3: the link itself counts down, see pre3);
// kPopLink: the entries looked at by the last pop were all culled, pop again; kDoneLink: finished / no ray.
//
// The pop is branch-free and bounded: it looks at the two topmost deferred siblings at once (two independent
// local-memory loads) and, if both are culled, leaves the lane in the kPopLink state for the next iteration.  ncu showed
// the `do {} while (m > cull)` form running at 2 active lanes with a quarter of all stall samples on its dependent
// loads (profiles/r01e_k_wf_trace_lines.txt).
template <bool COUNT, int RING = 0>
struct LaneTraversal {
    // the deferred-sibling stack lives in the caller's frame (dynamically indexed => local memory); keeping the arrays
    // out of this struct lets every other member be promoted to registers
    int *s_link;
    float *s_m;
    int2 *s2;               // form 3: one array of {link, entry parameter} pairs (one 64-bit local access per entry)
    int sp;
    int link;
    int tri_i, tri_end;
    float omax, cull;
    unsigned sgn;           // bit k: direction[k] > 0 (the near-child rule of bvh.cpp:183)
    Ray r;
    Hit best;

    RTB_HD void reset(int *stack_link, float *stack_m) {
        s_link = stack_link;
        s_m = stack_m;
        s_link[0] = kDoneLink;
        s_m[0] = -1.f;          // the sentinel is never culled
        sp = 1;
        link = kDoneLink;
        tri_i = tri_end = 0;
        omax = 0.f;
        cull = 0.f;
        sgn = 0u;
        best.t = kFar; best.u = 0.f; best.v = 0.f; best.id = -1;
        for (int k = 0; k < 3; ++k) { r.o[k] = 0.f; r.d[k] = 0.f; r.inv[k] = 0.f; }
    }
    RTB_HD void reset3(int2 *stack) {
        reset(reinterpret_cast<int *>(stack), reinterpret_cast<float *>(stack) + 1);   // (the first form's arrays are not used)
        s2 = stack;
        s2[0] = make_int2(kDoneLink, (int) f2u(-1.f));   // the sentinel is never culled
        s2[kUvSlot] = make_int2(0, 0);
        clear3();
    }
    // form 3 keeps what only the end of a ray needs out of the registers (the 48 / 56-register budgets of 10 / 9 CTAs
    // per SM): the barycentrics of the best hit live in a fixed local-memory slot above the stack (one predicated 64-bit
    // store per improvement, one load per ray) and max|origin| is recomputed when a hit improves.
    static constexpr int kUvSlot = kMaxStack + 2;          // (one entry of headroom above the deepest stack: the next chunk of a big leaf)
    static constexpr int kStack3 = kMaxStack + 3;          // entries of the form-3 array
    // The top kRing entries of the form-3 stack live in a ring (shared memory on the device, [slot][thread]: a 64-bit
    // access per lane is conflict-free whatever slot each lane is at, 2 wavefronts per warp instruction); the array in
    // local memory is the backing store: entries [0, sp - rn) are there, [sp - rn, sp) are in the ring, an entry is
    // written back to its place in the array only when a push finds the ring full.  Local memory interleaves 32-bit
    // words across the lanes, so a lane's 8-byte entry costs 2 sectors = 2 LSU wavefronts per LANE at divergent
    // depths: pushes and pops were 36 % of the LSU data-pipe work of the trace kernel, the unit that bounds it at 86 %
    // of peak (profiles/r01m_k_wf_trace_summary.txt).  kRing == 0: no ring, everything in the array.
    // RING == -1: no ring, but the TOP entry of the stack is kept in registers (`top`; entries [0, sp - 1) in the array):
    // a pop then needs one local load (the second entry) instead of two, and the entry that becomes the new top is
    // either that one or is fetched off the critical path.
    static constexpr bool kTop = RING < 0;
    static constexpr int kRing = RING > 0 ? RING : 0;
    static_assert((kRing & (kRing - 1)) == 0, "the ring size is a power of two");
    int rn;                                                // entries in the ring
    int2 top;                                              // kTop: entry sp - 1
#ifndef __CUDA_ARCH__
    int2 host_ring[kRing > 0 ? kRing : 1];
#endif
    RTB_HD int2 &ring_at(int i) {
#ifdef __CUDA_ARCH__
        __shared__ int2 ring[(kRing > 0 ? kRing : 1) * kRingThreads];
        return ring[(i & (kRing - 1)) * kRingThreads + threadIdx.x];
#else
        return host_ring[i & (kRing - 1)];
#endif
    }
    RTB_HD void clear3() {   // an empty stack: the sentinel only (entry 0 of the array always holds it)
        sp = 1;
        rn = 0;
        top = make_int2(kDoneLink, (int) f2u(-1.f));
    }
    RTB_HD void push3(const int2 e) {
        if (kTop) {
            s2[sp - 1] = top;
            top = e;
        } else if (kRing > 0) {
            int2 &slot = ring_at(sp);
            if (rn == kRing) s2[sp - kRing] = slot;       // the oldest entry makes room
            else ++rn;
            slot = e;
        } else {
            s2[sp] = e;
        }
        ++sp;
    }
    // entry i of the stack wherever it lives (suspending)
    RTB_HD int2 entry3(int i) { return (kTop && i == sp - 1) ? top : (kRing > 0 && i >= sp - rn) ? ring_at(i) : s2[i]; }
    RTB_HD void finish3() {
        const int2 uv = s2[kUvSlot];
        best.u = u2f((uint32_t) uv.x);
        best.v = u2f((uint32_t) uv.y);
    }
    // ---- suspending a ray (form 3): everything a later launch needs to carry on from exactly this state.  The ray
    // itself is re-made from the caller's ray record (make_ray is a pure function of it).  rec: kSuspHeader + sp entries.
    static constexpr int kSuspHeader = 5;
    RTB_HD void save3(int2 *rec) {
        rec[0] = make_int2(link, sp);
        rec[1] = make_int2(tri_i, tri_end);
        rec[2] = make_int2((int) f2u(best.t), best.id);
        rec[3] = make_int2((int) f2u(cull), 0);
        rec[4] = s2[kUvSlot];
        for (int i = 0; i < sp; ++i) rec[kSuspHeader + i] = entry3(i);
    }
    // `r` has been set (as for begin)
    RTB_HD void load3(const int2 *rec) {
        const int2 a = rec[0], b = rec[1], c = rec[2], d = rec[3];
        link = a.x; sp = a.y;
        tri_i = b.x; tri_end = b.y;
        best.t = u2f((uint32_t) c.x); best.id = c.y;
        cull = u2f((uint32_t) d.x);
        s2[kUvSlot] = rec[4];
        for (int i = 0; i < sp; ++i) s2[i] = rec[kSuspHeader + i];
        rn = 0;
        top = rec[kSuspHeader + sp - 1];
        omax = ray_omax(r);
        sgn = (r.d[0] > 0.f ? 1u : 0u) | (r.d[1] > 0.f ? 2u : 0u) | (r.d[2] > 0.f ? 4u : 0u);
    }
    RTB_HD bool wants_node() const { return link >= 0; }
    RTB_HD bool wants_tri() const { return tri_i < tri_end; }
    RTB_HD bool wants_pop() const { return link == kPopLink; }
    RTB_HD bool finished() const { return link == kDoneLink; }

    RTB_HD bool nonzero_dir() const { return r.d[0] != 0.f && r.d[1] != 0.f && r.d[2] != 0.f; }
    // leaf links decode without a branch: kDoneLink / kPopLink carry a zero count, i.e. an empty range; only leaves
    // with more than 7 triangles (count field 0, an index into the big-leaf table) take the rare branch
    RTB_HD void set_link(const DBvh &b, int l) {
        link = l;
        int count = (l >> 28) & 7;
        int first = l & 0x0FFFFFFF;
        if (l < 0 && count == 0 && first < 0x0FFFFFFE) {
            const int2 bl = ldg(&b.big_leaves[first]);
            first = bl.x;
            count = bl.y;
        }
        tri_i = l < 0 ? first : tri_i;
        tri_end = l < 0 ? first + count : tri_end;
    }
    RTB_HD void pop(const DBvh &b) {
        const int i1 = sp - 1, i2 = sp >= 2 ? sp - 2 : 0;
        const int l1 = s_link[i1], l2 = s_link[i2];
        const float m1 = s_m[i1], m2 = s_m[i2];
        const bool k1 = !(m1 > cull), k2 = !(m2 > cull);
        int l = k1 ? l1 : (k2 ? l2 : kPopLink);
        sp = k1 ? i1 : i2;
        if (l == kDoneLink) sp = 1;
        set_link(b, l);
    }
    // `r` has been set: root box test (bvh.cpp:240)
    template <class C>
    RTB_HD void begin(const DBvh &b, C *cnt) {
        best.t = kFar; best.u = 0.f; best.v = 0.f; best.id = -1;
        omax = ray_omax(r);
        cull = cull_bound(kFar, omax);
        sgn = (r.d[0] > 0.f ? 1u : 0u) | (r.d[1] > 0.f ? 2u : 0u) | (r.d[2] > 0.f ? 4u : 0u);
        float m0;
        if (COUNT) cnt->box += 1;
        if (box_test_t(r, b.root_box[0], b.root_box[1], b.root_box[2], b.root_box[3], b.root_box[4], b.root_box[5], m0)) set_link(b, b.root_link);
    }
    template <bool NZ, class C>
    RTB_HD void node_compute(const DBvh &b, const float4 na, const float4 nb, const float4 nc, const int4 nd, C *cnt) {
        float ml, mr;
        const bool hl = box_test_t<NZ>(r, na.x, na.y, na.z, na.w, nb.x, nb.y, ml);
        const bool hr = box_test_t<NZ>(r, nb.z, nb.w, nc.x, nc.y, nc.z, nc.w, mr);
        if (COUNT) cnt->box += 2;
        const bool left_first = ((sgn >> nd.z) & 1u) != 0u;   // direction[split_dim] > 0 (bvh.cpp:183)
        const int lA = left_first ? nd.x : nd.y, lB = left_first ? nd.y : nd.x;
        const float mA = left_first ? ml : mr, mB = left_first ? mr : ml;
        const bool vA = (left_first ? hl : hr) && !(mA > cull);
        const bool vB = (left_first ? hr : hl) && !(mB > cull);
        if (vA && vB) {
            s_link[sp] = lB;
            s_m[sp] = mB;
            ++sp;
            set_link(b, lA);
        } else if (vA || vB) {
            set_link(b, vA ? lA : lB);
        } else {
            pop(b);
        }
    }
    template <class C>
    RTB_HD void tri_compute(const DBvh &b, const float4 q0, const float4 q1, const float4 q2, C *cnt) {
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
        if (tri_i == tri_end) pop(b);
    }
    template <class C>
    RTB_HD void node_step(const DBvh &b, C *cnt) {
        const DNode *np = b.nodes + link;
        const float4 na = ldg(&np->a), nb = ldg(&np->b), nc = ldg(&np->c);
        const int4 nd = ldg(&np->d);
        node_compute<false>(b, na, nb, nc, nd, cnt);
    }
    template <class C>
    RTB_HD void tri_step(const DBvh &b, C *cnt) {
        const float4 *tp = b.tris + (size_t) tri_i * kTriStride;
        const float4 q0 = ldg(tp), q1 = ldg(tp + 1), q2 = ldg(tp + 2);
        tri_compute(b, q0, q1, q2, cnt);
    }
    // ---- form 3 (MODE 3 of trace_persistent): ONE pop site and ONE leaf-decode site per step, both at the top of the
    // step where the lanes are converged.  node_compute3 / tri_compute3 only leave `link` behind (an internal node, a
    // not yet decoded leaf, or kPopLink); pre3 pops for the lanes that asked for it and decodes fresh leaves.  In the
    // first form these two pieces were inlined at three places each and ran at 3 to 6 active lanes: 20 % of all
    // issued instructions (profiles/r01h_k_wf_trace_lines.txt).  Same visits, same arithmetic, same results.
    // In form 3 a leaf being worked on IS its link: {1, triangles left (3 bits), next triangle (28 bits)} -- no separate
    // range registers, nothing to decode; testing a triangle turns (n, i) into (n - 1, i + 1) and the last one into
    // kPopLink.  (unsigned) link >= kLeafBusy  <=>  a leaf link with at least one triangle left (kPopLink / kDoneLink
    // and big-leaf table links carry a zero count and lie below).
    static constexpr unsigned kLeafBusy = 0x90000000u;
    RTB_HD bool has_tri() const { return (unsigned) link >= kLeafBusy; }
    RTB_HD void pre3(const DBvh &b) {
        if (link == kPopLink) {
            const int i1 = sp - 1, i2 = sp >= 2 ? sp - 2 : 0;
            int2 e1, e2;
            if (kTop) {
                e1 = top;
                e2 = s2[i2];
            } else if (kRing > 0) {
                e1 = ring_at(i1);
                e2 = ring_at(i2);
                if (rn < 1) e1 = s2[i1];
                if (rn < 2) e2 = s2[i2];
            } else {
                e1 = s2[i1];
                e2 = s2[i2];     // (reading it only when e1 is culled saves sectors but chains two loads: -0.8 %, r01m_stack_ring_experiment.txt)
            }
            const bool k1 = !(u2f((uint32_t) e1.y) > cull), k2 = !(u2f((uint32_t) e2.y) > cull);
            link = k1 ? e1.x : (k2 ? e2.x : kPopLink);
            sp = k1 ? i1 : i2;
            if (kRing > 0) { rn -= k1 ? 1 : 2; rn = rn < 0 ? 0 : rn; }
            if (kTop) {
                top = e2;                                   // k1: the second entry is the new top
                if (!k1) top = s2[i2 >= 1 ? i2 - 1 : 0];    // both consumed: the one below (not needed before the next pop / push)
            }
            if (link == kDoneLink) sp = 1;                 // (the sentinel is entry 0 of the array, never in the ring)
        }
        if ((unsigned) link - 0x80000000u < 0x0FFFFFFEu) {   // a leaf of more than 7 triangles (rare): 7 now, the rest deferred
            const int2 bl = ldg(&b.big_leaves[link & 0x0FFFFFFF]);
            int take = bl.y;
            if (take > 7) {
                push3(make_int2(link + 1, (int) f2u(-1.f)));     // the next entry of the run; never culled: the leaf has been entered
                take = 7;
            }
            link = (int) (0x80000000u | ((unsigned) take << 28) | (unsigned) bl.x);
        }
    }
    template <class C>
    RTB_HD void node_compute3(const float4 na, const float4 nb, const float4 nc, const int4 nd, C *cnt) {
        float ml, mr;
        const bool hl = box_test_z(r, na.x, na.y, na.z, na.w, nb.x, nb.y, ml);
        const bool hr = box_test_z(r, nb.z, nb.w, nc.x, nc.y, nc.z, nc.w, mr);
        if (COUNT) cnt->box += 2;
        const bool vl = hl && !(ml > cull), vr = hr && !(mr > cull);
        const bool left_first = ((sgn >> nd.z) & 1u) != 0u;   // direction[split_dim] > 0 (bvh.cpp:183)
        const bool both = vl && vr;
        const bool take_left = vl && (left_first || !vr);    // both ? left_first : vl
        if (both) push3(make_int2(take_left ? nd.y : nd.x, (int) f2u(take_left ? mr : ml)));
        link = (vl || vr) ? (take_left ? nd.x : nd.y) : kPopLink;
    }
    template <class C>
    RTB_HD void tri_compute3(const float4 q0, const float4 q1, const float4 q2, C *cnt) {
        float t, u, v;
        if (COUNT) cnt->tri += 1;
        if (tri_test_flat(r, q0, q1, q2, t, u, v) && t < best.t) {
            best.t = t;
            best.id = link & 0x0FFFFFFF;
            s2[kUvSlot] = make_int2((int) f2u(u), (int) f2u(v));
            cull = cull_bound(t, ray_omax(r));
        }
        const int next = link - 0x0FFFFFFF;       // one triangle fewer, the next index
        link = (unsigned) next >= kLeafBusy ? next : kPopLink;
    }
    template <class C>
    RTB_HD void step3(const DBvh &b, C *cnt) {
        pre3(b);
        if (link >= 0) {
            const DNode *np = b.nodes + link;
            node_compute3(ldg(&np->a), ldg(&np->b), ldg(&np->c), ldg(&np->d), cnt);
        } else if (has_tri()) {
            const float4 *tp = b.tris + (size_t) (link & 0x0FFFFFFF) * kTriStride;
            tri_compute3(ldg(tp), ldg(tp + 1), ldg(tp + 2), cnt);
        }
    }
    // whatever the lane's state asks for (host drivers; the device loops below schedule the steps per warp)
    template <class C>
    RTB_HD void step(const DBvh &b, C *cnt) {
        if (wants_pop()) pop(b);
        else if (wants_node()) node_step(b, cnt);
        else if (wants_tri()) tri_step(b, cnt);
    }
};

#ifdef __CUDACC__
__device__ __forceinline__ void prefetch_l1(const void *p) { asm volatile("prefetch.global.L1 [%0];" ::"l"(p)); }
__device__ __forceinline__ void prefetch_l2(const void *p) { asm volatile("prefetch.global.L2 [%0];" ::"l"(p)); }
// 32 bytes through the read-only path in one instruction (p 32-byte aligned)
__device__ __forceinline__ void ldg256(const float4 *p, float4 &lo, float4 &hi) {
    asm volatile("ld.global.nc.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=f"(lo.x), "=f"(lo.y), "=f"(lo.z), "=f"(lo.w), "=f"(hi.x), "=f"(hi.y), "=f"(hi.z), "=f"(hi.w)
                 : "l"(p));
}
__device__ __forceinline__ void ldg256(const float4 *p, float4 &lo, int4 &hi) {
    asm volatile("ld.global.nc.v8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=f"(lo.x), "=f"(lo.y), "=f"(lo.z), "=f"(lo.w), "=r"(hi.x), "=r"(hi.y), "=r"(hi.z), "=r"(hi.w)
                 : "l"(p));
}
__device__ __forceinline__ void ldg256_noalloc(const float4 *p, float4 &lo, float4 &hi) {
    asm volatile("ld.global.nc.L1::no_allocate.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=f"(lo.x), "=f"(lo.y), "=f"(lo.z), "=f"(lo.w), "=f"(hi.x), "=f"(hi.y), "=f"(hi.z), "=f"(hi.w)
                 : "l"(p));
}
__device__ __forceinline__ void ldg256_noalloc(const float4 *p, float4 &lo, int4 &hi) {
    asm volatile("ld.global.nc.L1::no_allocate.v8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=f"(lo.x), "=f"(lo.y), "=f"(lo.z), "=f"(lo.w), "=r"(hi.x), "=r"(hi.y), "=r"(hi.z), "=r"(hi.w)
                 : "l"(p));
}
__device__ __forceinline__ void ldg256_last(const float4 *p, float4 &lo, float4 &hi) {
    asm volatile("ld.global.nc.L1::evict_last.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=f"(lo.x), "=f"(lo.y), "=f"(lo.z), "=f"(lo.w), "=f"(hi.x), "=f"(hi.y), "=f"(hi.z), "=f"(hi.w)
                 : "l"(p));
}
__device__ __forceinline__ void ldg256_last(const float4 *p, float4 &lo, int4 &hi) {
    asm volatile("ld.global.nc.L1::evict_last.v8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=f"(lo.x), "=f"(lo.y), "=f"(lo.z), "=f"(lo.w), "=r"(hi.x), "=r"(hi.y), "=r"(hi.z), "=r"(hi.w)
                 : "l"(p));
}
// The persistent loop of the queue kernels ("form 3", see LaneTraversal::pre3); MODE - 2 steps are taken between two
// refill votes (MODE 3 / 4 / 5).  Earlier forms -- one phase per warp iteration chosen by vote; the first unified step
// with the pop and the leaf decode inlined at three sites -- and what was tried and dropped (parked leaves, batching or
// gating the triangle code, a second instantiation for rays without zero direction components, consolidating the tail
// through an in-launch requeue) are in DESIGN.md section 5 with their ncu evidence.  Emit(tag, ray, hit) stores a result.
//
// Fetch(idx, L, tag&) loads queue entry idx into L.r and returns 0 (the entry holds no ray), 1 (a new ray: L.begin
// follows) or 2 (a ray suspended by an earlier launch: Fetch has restored the traversal state with L.load3).
// Suspend(tag, L): see `tail_steps`.
//
// tail_steps > 0 (form 3 only) bounds the TAIL of a launch: a ray that has run tail_steps steps in this launch while the
// queue is already empty is suspended -- Suspend() saves its traversal state (L.save3) and the lane retires -- and a
// later launch resumes it where it stopped.  Without this a launch lasts as long as its longest ray (hundreds of
// steps against a mean of ~50), which is what bounds small frames / many GPUs (profiles/r01k_*): the iterations of a
// pixel are a dependent chain.  The ray's visits, arithmetic and result are unchanged; only when they happen moves.
template <bool COUNT, int MODE, int RING, class Fetch, class Emit, class Suspend>
__device__ __forceinline__ void trace_persistent(const DBvh &b, unsigned n, unsigned *queue_head, Fetch fetch, Emit emit, Suspend suspend,
                                                 const int tail_steps, QueueCounters *cnt, const TracePolicy pol) {
    const unsigned FULL = 0xffffffffu;
    const int lane = threadIdx.x & 31;
    __align__(8) int2 stack[LaneTraversal<COUNT, RING>::kStack3];
    LaneTraversal<COUNT, RING> L;
    L.reset3(stack);
    unsigned tag = 0;
    bool have_ray = false, exhausted = false;
    int budget = tail_steps;

    while (true) {
        const bool wantR = L.finished() && !exhausted;
        const bool busy = !L.finished();
        const unsigned mR = __ballot_sync(FULL, wantR), mB = __ballot_sync(FULL, busy);
        if ((mR | mB) == 0u) break;
        const int cR = __popc(mR);
        if (cR >= pol.refill_min || mB == 0u) {
            // ---- refill: store finished results, claim new rays (warp-aggregated atomic)
            const int leader = __ffs(mR) - 1;
            unsigned base = 0;
            if (lane == leader) base = atomicAdd(queue_head, (unsigned) cR);
            base = __shfl_sync(FULL, base, leader);
            if (wantR) {
                if (have_ray) {
                    L.finish3();
                    emit(tag, L.r, L.best);
                }
                const unsigned idx = base + (unsigned) __popc(mR & ((1u << lane) - 1u));
                if (idx < n) {
                    const int got = fetch(idx, L, tag);
                    have_ray = got != 0;
                    if (got == 1) {
                        cnt->rays += 1;
                        L.begin(b, cnt);
                    }
                    budget = tail_steps;
                } else {
                    have_ray = false;
                    exhausted = true;
                }
            }
            continue;
        }
        // MODE 3 / 4 / 5: one / two / three steps per refill vote (profiles/r01j_sched_sweep.txt: two is best, +2 %).
        // Gating the triangle code on the number of lanes holding a triangle (they wait a step unless >= 4 / 8 / 12
        // want it) was measured there too and loses 4 / 6 / 10 %: a ray's latency matters more than the issue slots.
#pragma unroll
        for (int step = 0; step < MODE - 2; ++step) {
            L.pre3(b);
            const bool doN = L.link >= 0;
            const bool doL = L.has_tri();
            // One 64-byte record per lane -- its node or its triangle -- with two 256-bit loads (LDG.E.256, new on
            // sm_100): half the load instructions and half the LSU wavefronts of the 4 x LDG.128 form, which ncu showed
            // at 81 % of the LSU data-pipe peak (profiles/r01i_k_wf_trace_summary.txt).  Idle lanes (waiting for a
            // refill or for their next pop) read triangle 0: no predicates, no zeroing.
            const float4 *p = doN ? reinterpret_cast<const float4 *>(b.nodes + L.link) : (b.tris + (size_t) (doL ? L.link & 0x0FFFFFFF : 0) * kTriStride);
            float4 a, bb, c;
            int4 d;
#if RTB_FETCH_HINTS
            if (doN) {
#if RTB_FETCH_HINTS >= 2
                ldg256_last(p, a, bb);
                ldg256_last(p + 2, c, d);
#else
                ldg256(p, a, bb);
                ldg256(p + 2, c, d);
#endif
            } else {
                ldg256_noalloc(p, a, bb);
                ldg256_noalloc(p + 2, c, d);
            }
#else
            ldg256(p, a, bb);
            ldg256(p + 2, c, d);
#endif
#if RTB_PREFETCH_LEAF
            // the next triangle of the leaf (if any) will be fetched by this lane's next step whatever this test says
            if (doL && (unsigned) (L.link - 0x0FFFFFFF) >= LaneTraversal<COUNT, RING>::kLeafBusy) prefetch_l1(p + kTriStride);
#endif
#if RTB_PREFETCH_SIBLING
            const int sp_before = L.sp;
#endif
            if (doN) L.node_compute3(a, bb, c, d, cnt);
#if RTB_PREFETCH_SIBLING
            if (doN && L.sp > sp_before) {   // a sibling was deferred: it is needed when this subtree is done
                const int dl = L.entry3(L.sp - 1).x;
                prefetch_l1(dl >= 0 ? reinterpret_cast<const float4 *>(b.nodes + dl) : b.tris + (size_t) (dl & 0x0FFFFFFF) * kTriStride);
            }
#endif
            if (doL) L.tri_compute3(a, bb, c, cnt);
        }
        if (tail_steps > 0) {
            budget -= MODE - 2;
            if (budget <= 0 && !L.finished()) {
                // this ray has had its steps: if the queue has run dry the launch is in its tail -- step aside
                if (*reinterpret_cast<volatile unsigned *>(queue_head) >= n) {
                    suspend(tag, L);
                    have_ray = false;
                    L.link = kDoneLink;
                    L.clear3();
                } else {
                    budget = 16;
                }
            }
        }
    }
}
#endif

}  // namespace rtb
