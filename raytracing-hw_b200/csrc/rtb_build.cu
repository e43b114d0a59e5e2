// GPU BVH builder (sm_100a) with the reference's node AND triangle layout: BVH::buildBVH (src/core/bvh.cpp:5-175).
//
// The reference builds top-down: at every node it std::sorts the node's triangles by the centroid on the node's widest
// axis (bvh.cpp:47-50), sweeps the SAH over every cut of the sorted range (bvh.cpp:53-92), cuts where it is cheapest and
// std::partitions (bvh.cpp:109-110).  The layout is what libstdc++'s introsort leaves behind, ties included, so the sort
// is not replaced by "a" sort: it is REPLAYED (csrc/rtb_build_steps.cuh: where introsort's partition cursors stop is a
// function of the range before the first swap, hence two prefix sums, a scatter and an exchange; the recursion runs
// breadth first, one partition step of every live range per round; the blocks of <= 16 it leaves are insertion-sorted,
// ranges that run out of depth are heap-sorted with libstdc++'s own heap moves).
//
// Everything is level-synchronous over the WHOLE triangle array: a level of the tree = all nodes of that depth that
// still have at least `host_below` triangles, processed by the same passes --
//   boxes of the nodes (segmented in-order scan: AABB::grow keeps the earlier of two equal bounds)  ->  widest axis,
//   keys  ->  the sort rounds  ->  prefix / suffix box scans in sorted order (segmented scans with the reference's
//   grow order)  ->  SAH metric of every cut, first minimum per node (64-bit atomicMin of {metric bits, cut})  ->
//   cut, children.
// HBM-bound byte work (8..28 bytes per triangle and pass, coalesced); no tensor cores, nothing to contract.
// Nodes the kernels do not cut (fewer than host_below triangles; large nodes without any SAH improvement, whose
// std::partition on z is replayed on the host) are returned as ranges of the order array for the host builder.
#include "rtb_build.h"

#ifdef RTB_BUILD_EMULATE   // tests/devsim/build_emul: this file compiled for the host, kernels run as loops (test tooling)
#include "cuda_emulation.h"
#else
#include <cuda_runtime.h>
#endif

#include <algorithm>
#include <chrono>
#include <cstdio>

#include "rtb_build_steps.cuh"

namespace rtb {
namespace {
using namespace gbuild;

constexpr int kScanThreads = 256, kScanItems = 4, kScanTile = kScanThreads * kScanItems;
constexpr uint32_t kSerialPartitionMax = 256;   // a node without SAH improvement up to this size replays std::partition in its thread

struct SegBox {
    Box6 b;
    uint32_t head;     // first element of its segment (in scan order)
};
struct SegBoxOp {
    __device__ __forceinline__ SegBox operator()(const SegBox &a, const SegBox &b) const {
        if (b.head) return b;
        SegBox r;
        r.b = box_grow(a.b, b.b);
        r.head = a.head;
        return r;
    }
};
struct SumOp {
    __device__ __forceinline__ unsigned long long operator()(unsigned long long a, unsigned long long b) const { return a + b; }
};

#ifndef RTB_BUILD_EMULATE
// ---- inclusive scan with an associative (not necessarily commutative) operator, three passes
template <class T, class Op>
__global__ void __launch_bounds__(kScanThreads) k_scan_tiles(const T *in, T *out, T *__restrict__ agg, uint32_t n, Op op) {
    __shared__ T s[kScanThreads];
    const uint32_t tile = blockIdx.x * (uint32_t) kScanTile;
    const uint32_t base = tile + threadIdx.x * kScanItems;
    T v[kScanItems];
    int m = 0;
#pragma unroll
    for (int k = 0; k < kScanItems; ++k)
        if (base + k < n) { v[k] = in[base + k]; m = k + 1; }
#pragma unroll
    for (int k = 1; k < kScanItems; ++k)
        if (k < m) v[k] = op(v[k - 1], v[k]);
    if (m > 0) s[threadIdx.x] = v[m - 1];
    __syncthreads();
    for (int d = 1; d < kScanThreads; d <<= 1) {
        T t;
        const bool on = (int) threadIdx.x >= d && m > 0;
        if (on) t = op(s[threadIdx.x - d], s[threadIdx.x]);
        __syncthreads();
        if (on) s[threadIdx.x] = t;
        __syncthreads();
    }
    if (threadIdx.x > 0 && m > 0) {
        const T p = s[threadIdx.x - 1];
#pragma unroll
        for (int k = 0; k < kScanItems; ++k)
            if (k < m) v[k] = op(p, v[k]);
    }
#pragma unroll
    for (int k = 0; k < kScanItems; ++k)
        if (k < m) out[base + k] = v[k];
    const uint32_t left = n - tile < (uint32_t) kScanTile ? n - tile : (uint32_t) kScanTile;
    if (threadIdx.x == (left - 1) / kScanItems) agg[blockIdx.x] = s[threadIdx.x];
}
// one block: inclusive scan of the tile aggregates in place
template <class T, class Op>
__global__ void __launch_bounds__(kScanThreads) k_scan_aggs(T *__restrict__ agg, uint32_t n_tiles, Op op) {
    __shared__ T s[kScanThreads];
    __shared__ T carry;
    for (uint32_t base = 0; base < n_tiles; base += kScanThreads) {
        const uint32_t i = base + threadIdx.x;
        const bool have = i < n_tiles;
        if (have) s[threadIdx.x] = agg[i];
        __syncthreads();
        for (int d = 1; d < kScanThreads; d <<= 1) {
            T t;
            const bool on = (int) threadIdx.x >= d && have;
            if (on) t = op(s[threadIdx.x - d], s[threadIdx.x]);
            __syncthreads();
            if (on) s[threadIdx.x] = t;
            __syncthreads();
        }
        T r;
        if (have) r = base ? op(carry, s[threadIdx.x]) : s[threadIdx.x];
        __syncthreads();
        if (have) {
            agg[i] = r;
            if (threadIdx.x == kScanThreads - 1 || i == n_tiles - 1) carry = r;
        }
        __syncthreads();
    }
}
template <class T, class Op>
__global__ void __launch_bounds__(kScanThreads) k_scan_apply(T *__restrict__ out, const T *__restrict__ agg, uint32_t n, Op op) {
    if (blockIdx.x == 0) return;
    const T p = agg[blockIdx.x - 1];
    const uint32_t base = blockIdx.x * (uint32_t) kScanTile + threadIdx.x * kScanItems;
#pragma unroll
    for (int k = 0; k < kScanItems; ++k)
        if (base + k < n) out[base + k] = op(p, out[base + k]);
}
template <class T, class Op>
void scan_inclusive(const T *in, T *out, T *agg, uint32_t n, Op op, cudaStream_t st) {
    if (n == 0) return;
    const uint32_t tiles = (n + kScanTile - 1) / kScanTile;
    k_scan_tiles<T, Op><<<tiles, kScanThreads, 0, st>>>(in, out, agg, n, op);
    if (tiles > 1) {
        k_scan_aggs<T, Op><<<1, kScanThreads, 0, st>>>(agg, tiles, op);
        k_scan_apply<T, Op><<<tiles, kScanThreads, 0, st>>>(out, agg, n, op);
    }
}

#else
// (host emulation: threads of a block run one after the other, so the block-wide scans above cannot; same operator, in order)
template <class T, class Op>
void scan_inclusive(const T *in, T *out, T *, uint32_t n, Op op, cudaStream_t) {
    if (n == 0) return;
    T acc = in[0];
    out[0] = acc;
    for (uint32_t i = 1; i < n; ++i) {
        acc = op(acc, in[i]);
        out[i] = acc;
    }
}
#endif

// ---- a level of the tree
struct LevelNode {
    uint32_t first, count;
    uint32_t top;              // slot of the node in the output array
    int32_t dim;               // widest axis
    Box6 box;
    float first_z;             // centre z of the range's first triangle before the sort (bvh.cpp:20)
    uint32_t lc;               // triangles left of the cut
    unsigned long long best;   // {metric bits, cut} of the cheapest cut
    uint32_t child[2];         // the children's indices in the next level (kNoRange: not cut further here)
    int32_t restore;           // the range goes back to its order before the sort (handed to the host uncut)
    uint32_t pad;
};

struct Counters {
    uint32_t n_ranges, active, heaps, nn_next, n_top, fallback, overflow, pad;
};

struct Dev {
    uint32_t n;
    const Box6 *tbox;
    const float *ctr;
    uint32_t *idx, *idx_prev, *rid, *nid, *nid_next;
    float *key;
    unsigned long long *flags, *scan;
    uint32_t *posL, *posR;
    SegBox *F, *Fs, *R, *Rs;
    SortRange *ranges;
    SortRound *rr;
    LevelNode *nodes, *next;
    GpuTopNode *top;
    Counters *cnt;
    uint32_t top_cap, host_below;
    int depth_override;
};

__global__ void k_iota(uint32_t *idx, uint32_t *nid, uint32_t n) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) { idx[i] = i; nid[i] = 0u; }
}

// boxes of the level's nodes: the triangles' boxes in array order, a segment per node (elements outside: their own segment)
__global__ void k_seg_boxes(const Dev d) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= d.n) return;
    const uint32_t nd = d.nid[i];
    SegBox s;
    s.b = d.tbox[d.idx[i]];
    s.head = nd == kNoRange ? 1u : (i == d.nodes[nd].first ? 1u : 0u);
    d.F[i] = s;
}

__global__ void k_node_setup(const Dev d, uint32_t nn) {
    const uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= nn) return;
    LevelNode nd = d.nodes[j];
    nd.box = d.Fs[nd.first + nd.count - 1].b;
    nd.dim = box_widest(nd.box);
    nd.first_z = d.ctr[(size_t) d.idx[nd.first] * 3 + 2];
    nd.best = ~0ull;
    nd.lc = 0;
    nd.child[0] = nd.child[1] = kNoRange;
    nd.restore = 0;
    d.nodes[j] = nd;
    const int depth = d.depth_override >= 0 ? d.depth_override : 2 * lg_floor(nd.count);
    SortRange r = {nd.first, nd.first + nd.count, depth, range_state(nd.first, nd.first + nd.count, depth)};
    d.ranges[j] = r;
    if (r.state == kRangeActive) atomicAdd(&d.cnt->active, 1u);
    if (r.state == kRangeHeap) atomicAdd(&d.cnt->heaps, 1u);
}

__global__ void k_keys(const Dev d) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= d.n) return;
    const uint32_t nd = d.nid[i];
    const uint32_t t = d.idx[i];
    d.idx_prev[i] = t;
    d.rid[i] = nd;
    if (nd != kNoRange) d.key[i] = d.ctr[(size_t) t * 3 + d.nodes[nd].dim];
}

// ---- one round of the sort: see csrc/rtb_build_steps.cuh and tests/devsim/build_check.cpp (the same passes as loops)
__global__ void k_pivot(const Dev d, uint32_t n_ranges) {
    const uint32_t r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= n_ranges) return;
    SortRound q = {0.f, 0u, kNoRange, 0u};
    const SortRange g = d.ranges[r];
    if (g.state == kRangeActive) {
        median_to_first(d.key, d.idx, g.lo, g.hi);
        q.pivot = d.key[g.lo];
    }
    d.rr[r] = q;
}
__global__ void k_flags(const Dev d) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= d.n) return;
    unsigned long long f = 0ull;
    const uint32_t r = d.rid[i];
    if (r != kNoRange) {
        const SortRange g = d.ranges[r];
        if (g.state == kRangeActive && i > g.lo) {
            const float p = d.rr[r].pivot, k = d.key[i];
            if (stops_left(k, p)) f |= 1ull;
            if (stops_right(k, p)) f |= 1ull << 32;
        }
    }
    d.flags[i] = f;
}
__global__ void k_positions(const Dev d) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= d.n) return;
    const unsigned long long f = d.flags[i];
    if (!f) return;
    const SortRange g = d.ranges[d.rid[i]];
    const unsigned long long base = d.scan[g.lo], end = d.scan[g.hi - 1], me = d.scan[i];
    if (f & 1ull) d.posL[g.lo + 1u + ((uint32_t) me - (uint32_t) base - 1u)] = i;
    if (f >> 32) d.posR[g.lo + 1u + ((uint32_t) (end >> 32) - (uint32_t) (me >> 32))] = i;
}
__global__ void k_exchange(const Dev d) {
    const uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= d.n) return;
    const uint32_t r = d.rid[j];
    if (r == kNoRange) return;
    const SortRange g = d.ranges[r];
    if (g.state != kRangeActive || j <= g.lo) return;
    const uint32_t k = j - (g.lo + 1u);
    const unsigned long long base = d.scan[g.lo], end = d.scan[g.hi - 1];
    const uint32_t nL = (uint32_t) end - (uint32_t) base;
    const uint32_t nR = (uint32_t) (end >> 32) - (uint32_t) (base >> 32);
    uint32_t a = 0, b = 0;
    const bool v = k < nL && k < nR && (a = d.posL[j]) < (b = d.posR[j]);
    if (k == 0u && !v) d.rr[r].cut = d.posL[g.lo + 1u];
    if (v) {
        const bool more = k + 1u < nL && k + 1u < nR && d.posL[j + 1u] < d.posR[j + 1u];
        if (!more) {
            const uint32_t l_next = k + 1u < nL ? d.posL[j + 1u] : kInf;
            d.rr[r].cut = l_next < b ? l_next : b;
        }
        swap_at(d.key, d.idx, a, b);
    }
}
__global__ void k_split(const Dev d, uint32_t n_ranges) {
    const uint32_t r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= n_ranges) return;
    SortRange g = d.ranges[r];
    if (g.state != kRangeActive) return;
    const uint32_t cut = d.rr[r].cut;
    const int dep = g.depth - 1;
    const uint32_t id = atomicAdd(&d.cnt->n_ranges, 1u);
    SortRange right = {cut, g.hi, dep, range_state(cut, g.hi, dep)};
    d.ranges[id] = right;
    d.rr[r].right_id = id;
    g.hi = cut;
    g.depth = dep;
    g.state = range_state(g.lo, cut, dep);
    d.ranges[r] = g;
    const uint32_t act = (right.state == kRangeActive ? 1u : 0u) + (g.state == kRangeActive ? 1u : 0u);
    if (act) atomicAdd(&d.cnt->active, act);
    const uint32_t hp = (right.state == kRangeHeap ? 1u : 0u) + (g.state == kRangeHeap ? 1u : 0u);
    if (hp) atomicAdd(&d.cnt->heaps, hp);
}
__global__ void k_assign(const Dev d, uint32_t n_ranges) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= d.n) return;
    const uint32_t r = d.rid[i];
    if (r == kNoRange || r >= n_ranges) return;
    const SortRound q = d.rr[r];
    if (q.right_id != kNoRange && i >= q.cut) d.rid[i] = q.right_id;
}
__global__ void k_finish(const Dev d, uint32_t n_ranges) {
    const uint32_t r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= n_ranges) return;
    const SortRange g = d.ranges[r];
    if (g.state == kRangeSmall) insertion_sort_block(d.key, d.idx, g.lo, g.hi);
    else if (g.state == kRangeHeap) partial_sort_exact(d.key, d.idx, g.lo, g.hi);
}

// ---- SAH sweep over the sorted ranges
__global__ void k_sorted_boxes(const Dev d) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= d.n) return;
    const uint32_t nd = d.nid[i];
    SegBox s;
    s.b = d.tbox[d.idx[i]];
    uint32_t hf = 1u, hb = 1u;
    if (nd != kNoRange) {
        const uint32_t first = d.nodes[nd].first, count = d.nodes[nd].count;
        hf = i == first ? 1u : 0u;
        hb = i == first + count - 1u ? 1u : 0u;
    }
    s.head = hf;
    d.F[i] = s;
    s.head = hb;
    d.R[d.n - 1u - i] = s;     // the right scan runs from the last triangle down (bvh.cpp:70-75)
}
__global__ void k_metric(const Dev d) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    const uint32_t nd = i < d.n ? d.nid[i] : kNoRange;
    unsigned long long packed = ~0ull;
    if (nd != kNoRange) {
        const uint32_t first = d.nodes[nd].first, count = d.nodes[nd].count;
        if (i + 1u < first + count) {
            const uint32_t left = i - first + 1u;
            const float ls = box_area(d.Fs[i].b);
            const float rs = box_area(d.Rs[d.n - 1u - (i + 1u)].b);
            packed = pack_metric(sah_metric(ls, rs, left, count), left);
        }
    }
    // one atomic per warp when the warp sits inside one node (the common case), else one per lane
    const uint32_t nd0 = __shfl_sync(0xffffffffu, nd, 0);
    if (__all_sync(0xffffffffu, nd == nd0)) {
        for (int o = 16; o > 0; o >>= 1) {
            const unsigned long long other = __shfl_down_sync(0xffffffffu, packed, o);
            packed = other < packed ? other : packed;
        }
        if ((threadIdx.x & 31) == 0 && nd != kNoRange && packed != ~0ull) atomicMin(&d.nodes[nd].best, packed);
    } else if (packed != ~0ull) {
        atomicMin(&d.nodes[nd].best, packed);
    }
}
__global__ void k_decide(const Dev d, uint32_t nn) {
    const uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= nn) return;
    LevelNode nd = d.nodes[j];
    const uint32_t first = nd.first, count = nd.count;
    GpuTopNode t;
    for (int k = 0; k < 3; ++k) { t.box[k] = nd.box.mn[k]; t.box[3 + k] = nd.box.mx[k]; }
    t.left = t.right = -1;
    t.first = first;
    t.count = count;
    const float best_sah = box_area(nd.box) * (float) count;   // bvh.cpp:19
    union { float f; uint32_t u; } mb;
    mb.u = (uint32_t) (nd.best >> 32);
    uint32_t lc = 0;
    bool fallback = false;
    if (nd.best != ~0ull && mb.f < best_sah) {
        // the first cheapest cut; the cut value is the mean of the two centroids either side of it (bvh.cpp:85-91), and
        // std::partition on a range sorted by the same key moves nothing: its result is the number of keys below it
        const uint32_t res = (uint32_t) nd.best;
        const float best_split = (d.key[first + res - 1u] + d.key[first + res]) * 0.5f;
        uint32_t lo = 0, hi = count;
        while (lo < hi) {
            const uint32_t mid = lo + (hi - lo) / 2u;
            if (d.key[first + mid] < best_split) lo = mid + 1u;
            else hi = mid;
        }
        lc = lo;
        t.split_dim = nd.dim;
    } else if (count <= kSerialPartitionMax) {
        // no cut beats the node's own cost: best_dim stays 2 and best_split the first triangle's centre z minus 1e-7
        // (bvh.cpp:20-22); std::partition (bidirectional, bits/stl_algo.h) on the sorted range with that predicate
        const float bs = (float) ((double) nd.first_z - 1e-7);
        uint32_t lo = first, hi = first + count;
        while (true) {
            while (true) {
                if (lo == hi) goto done;
                if (d.ctr[(size_t) d.idx[lo] * 3 + 2] < bs) ++lo;
                else break;
            }
            --hi;
            while (true) {
                if (lo == hi) goto done;
                if (!(d.ctr[(size_t) d.idx[hi] * 3 + 2] < bs)) --hi;
                else break;
            }
            swap_at(d.key, d.idx, lo, hi);
            ++lo;
        }
    done:
        lc = lo - first;
        t.split_dim = 2;
    } else {
        fallback = true;
    }
    if (fallback) {
        t.kind = kTopHost;
        t.split_dim = -1;
        nd.restore = 1;
        atomicAdd(&d.cnt->fallback, 1u);
    } else if (lc == 0u || lc == count) {
        t.kind = kTopLeaf;     // bvh.cpp:143-147
    } else {
        const uint32_t pair = atomicAdd(&d.cnt->n_top, 2u);
        if (pair + 2u > d.top_cap) {
            atomicAdd(&d.cnt->overflow, 1u);
            t.kind = kTopLeaf;
        } else {
            t.kind = kTopInternal;
            t.left = (int32_t) pair;
            t.right = (int32_t) pair + 1;
            nd.lc = lc;
            for (int c = 0; c < 2; ++c) {
                const uint32_t cf = c ? first + lc : first, cc = c ? count - lc : lc;
                if (cc >= d.host_below) {
                    const uint32_t slot = atomicAdd(&d.cnt->nn_next, 1u);
                    LevelNode ch;
                    ch.first = cf;
                    ch.count = cc;
                    ch.top = pair + (uint32_t) c;
                    d.next[slot] = ch;
                    nd.child[c] = slot;
                } else {
                    GpuTopNode h;
                    for (int k = 0; k < 6; ++k) h.box[k] = 0.f;
                    h.left = h.right = -1;
                    h.split_dim = -1;
                    h.kind = kTopHost;
                    h.first = cf;
                    h.count = cc;
                    if (cc <= 4u) {
                        // a leaf of the reference's tree (bvh.cpp:134-139): its box is the in-order union of its triangles' boxes
                        // (bvh.cpp:130-132; growing the first box from itself changes nothing)
                        Box6 b = d.tbox[d.idx[cf]];
                        for (uint32_t i = 1; i < cc; ++i) b = box_grow(b, d.tbox[d.idx[cf + i]]);
                        for (int k = 0; k < 3; ++k) { h.box[k] = b.mn[k]; h.box[3 + k] = b.mx[k]; }
                        h.kind = kTopLeaf;
                    }
                    d.top[pair + (uint32_t) c] = h;
                }
            }
        }
    }
    d.top[nd.top] = t;
    d.nodes[j] = nd;
}
// the elements' nodes in the next level; ranges handed back uncut return to their order before the sort
__global__ void k_next_level(const Dev d) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= d.n) return;
    const uint32_t nd = d.nid[i];
    uint32_t nx = kNoRange;
    if (nd != kNoRange) {
        const LevelNode &p = d.nodes[nd];
        if (p.restore) d.idx[i] = d.idx_prev[i];
        else if (p.lc) nx = p.child[i < p.first + p.lc ? 0 : 1];
    }
    d.nid_next[i] = nx;
}

struct DeviceArena {
    std::vector<void *> ptrs;
    bool ok = true;
    template <class T>
    T *take(size_t count) {
        void *p = nullptr;
        if (!ok || cudaMalloc(&p, std::max<size_t>(count, 1) * sizeof(T)) != cudaSuccess) { ok = false; return nullptr; }
        ptrs.push_back(p);
        return (T *) p;
    }
    ~DeviceArena() {
        for (void *p : ptrs) cudaFree(p);
    }
};

inline unsigned grid_for(size_t n, int block) { return (unsigned) ((n + (size_t) block - 1) / (size_t) block); }

}  // namespace

bool gpu_build_top(const float *tbox, const float *tcenter, size_t n_in, size_t host_below, int device, int depth_override,
                   std::vector<GpuTopNode> &nodes_out, std::vector<uint32_t> &order, GpuBuildStats &stats, std::string &err) {
    const auto t_begin = std::chrono::steady_clock::now();
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev < 1) { cudaGetLastError(); err = "no CUDA device"; return false; }
    if (device < 0) { if (cudaGetDevice(&device) != cudaSuccess) { cudaGetLastError(); device = 0; } }   // the caller's current device
    if (device >= ndev) device = 0;
    if (n_in == 0 || n_in >= ((size_t) 1 << 30)) { err = "triangle count out of range for the GPU builder"; return false; }
    host_below = std::max<size_t>(5, host_below);
    int prev_device = 0;
    cudaGetDevice(&prev_device);
    struct Restore {
        int dev;
        ~Restore() { cudaSetDevice(dev); }
    } restore{prev_device};
    if (cudaSetDevice(device) != cudaSuccess) { cudaGetLastError(); err = "cudaSetDevice failed"; return false; }
    const uint32_t n = (uint32_t) n_in;

    DeviceArena A;
    Dev d{};
    d.n = n;
    d.host_below = (uint32_t) std::min<size_t>(host_below, 0x7fffffffu);
    d.depth_override = depth_override;
    Box6 *d_tbox = A.take<Box6>(n);
    float *d_ctr = A.take<float>((size_t) n * 3);
    d.tbox = d_tbox;
    d.ctr = d_ctr;
    d.idx = A.take<uint32_t>(n);
    d.idx_prev = A.take<uint32_t>(n);
    d.rid = A.take<uint32_t>(n);
    d.nid = A.take<uint32_t>(n);
    d.nid_next = A.take<uint32_t>(n);
    d.key = A.take<float>(n);
    // the box scans run in place (a tile reads its elements before it writes them); the sort's scratch -- 24 bytes per
    // triangle -- lives in the first box array (the two phases of a level never overlap)
    d.F = A.take<SegBox>((size_t) n + 1);
    d.Fs = d.F;
    d.R = A.take<SegBox>(n);
    d.Rs = d.R;
    d.flags = reinterpret_cast<unsigned long long *>(d.F);
    d.scan = d.flags + n;
    d.posL = reinterpret_cast<uint32_t *>(d.scan + n);
    d.posR = d.posL + n;
    const size_t tiles = ((size_t) n + kScanTile - 1) / kScanTile;
    SegBox *agg_box = A.take<SegBox>(tiles);
    unsigned long long *agg_sum = A.take<unsigned long long>(tiles);
    d.ranges = A.take<SortRange>((size_t) n + 1);
    d.rr = A.take<SortRound>((size_t) n + 1);
    const size_t node_cap = (size_t) n / host_below + 2;
    d.nodes = A.take<LevelNode>(node_cap);
    d.next = A.take<LevelNode>(node_cap);
    // a legal tree of n triangles has at most 2 n - 1 nodes (the SAH may peel single triangles off): room for all of them
    // (48 bytes each), so that an unbalanced scene does not throw the finished GPU build away
    const size_t top_cap = 2 * (size_t) n + 1;
    d.top_cap = (uint32_t) top_cap;
    d.top = A.take<GpuTopNode>(top_cap);
    d.cnt = A.take<Counters>(1);
    if (!A.ok) { cudaGetLastError(); err = "the GPU builder could not allocate its work arrays"; return false; }

    cudaStream_t st = nullptr;
    if (cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking) != cudaSuccess) { cudaGetLastError(); err = "cudaStreamCreate failed"; return false; }
    struct StreamGuard {
        cudaStream_t s;
        ~StreamGuard() { cudaStreamDestroy(s); }
    } guard{st};
    bool cuda_ok = true;
    auto chk = [&](cudaError_t e) { if (e != cudaSuccess && cuda_ok) { cuda_ok = false; err = std::string("GPU builder: ") + cudaGetErrorString(e); } };
    chk(cudaMemcpyAsync(d_tbox, tbox, (size_t) n * sizeof(Box6), cudaMemcpyHostToDevice, st));
    chk(cudaMemcpyAsync(d_ctr, tcenter, (size_t) n * 3 * sizeof(float), cudaMemcpyHostToDevice, st));
    const int B = 256;
    k_iota<<<grid_for(n, B), B, 0, st>>>(d.idx, d.nid, n);
    LevelNode root{};
    root.first = 0;
    root.count = n;
    root.top = 0;
    chk(cudaMemcpyAsync(d.nodes, &root, sizeof root, cudaMemcpyHostToDevice, st));
    Counters c{};
    c.n_top = 1;
    chk(cudaMemcpyAsync(d.cnt, &c, sizeof c, cudaMemcpyHostToDevice, st));
    chk(cudaStreamSynchronize(st));
    const auto t_kernels = std::chrono::steady_clock::now();

    uint32_t nn = 1;
    while (nn > 0 && cuda_ok) {
        ++stats.levels;
        // node boxes in array order, widest axes, keys
        k_seg_boxes<<<grid_for(n, B), B, 0, st>>>(d);
        scan_inclusive(d.F, d.Fs, agg_box, n, SegBoxOp(), st);
        c.n_ranges = nn;
        c.active = 0;
        c.nn_next = 0;
        chk(cudaMemcpyAsync(d.cnt, &c, 2 * sizeof(uint32_t), cudaMemcpyHostToDevice, st));                 // n_ranges, active
        chk(cudaMemcpyAsync(&d.cnt->nn_next, &c.nn_next, sizeof(uint32_t), cudaMemcpyHostToDevice, st));
        k_node_setup<<<grid_for(nn, B), B, 0, st>>>(d, nn);
        k_keys<<<grid_for(n, B), B, 0, st>>>(d);
        chk(cudaMemcpyAsync(&c, d.cnt, sizeof c, cudaMemcpyDeviceToHost, st));
        chk(cudaStreamSynchronize(st));
        // the sort: one partition step of every live range per round
        uint32_t n_ranges = nn;
        while (c.active > 0 && cuda_ok) {
            ++stats.rounds;
            const uint32_t zero = 0;
            chk(cudaMemcpyAsync(&d.cnt->active, &zero, sizeof zero, cudaMemcpyHostToDevice, st));
            k_pivot<<<grid_for(n_ranges, B), B, 0, st>>>(d, n_ranges);
            k_flags<<<grid_for(n, B), B, 0, st>>>(d);
            scan_inclusive(d.flags, d.scan, agg_sum, n, SumOp(), st);
            k_positions<<<grid_for(n, B), B, 0, st>>>(d);
            k_exchange<<<grid_for(n, B), B, 0, st>>>(d);
            k_split<<<grid_for(n_ranges, B), B, 0, st>>>(d, n_ranges);
            k_assign<<<grid_for(n, B), B, 0, st>>>(d, n_ranges);
            chk(cudaMemcpyAsync(&c, d.cnt, sizeof c, cudaMemcpyDeviceToHost, st));
            chk(cudaStreamSynchronize(st));
            n_ranges = c.n_ranges;
            if (n_ranges > n) { cuda_ok = false; err = "GPU builder: range table overflow"; }
        }
        if (!cuda_ok) break;
        k_finish<<<grid_for(n_ranges, 64), 64, 0, st>>>(d, n_ranges);
        // SAH sweep, cut, children
        k_sorted_boxes<<<grid_for(n, B), B, 0, st>>>(d);
        scan_inclusive(d.F, d.Fs, agg_box, n, SegBoxOp(), st);
        scan_inclusive(d.R, d.Rs, agg_box, n, SegBoxOp(), st);
        k_metric<<<grid_for(n, B), B, 0, st>>>(d);
        k_decide<<<grid_for(nn, 128), 128, 0, st>>>(d, nn);
        k_next_level<<<grid_for(n, B), B, 0, st>>>(d);
        chk(cudaMemcpyAsync(&c, d.cnt, sizeof c, cudaMemcpyDeviceToHost, st));
        chk(cudaStreamSynchronize(st));
        chk(cudaGetLastError());
        if (c.overflow) { cuda_ok = false; err = "GPU builder: node table overflow"; break; }
        std::swap(d.nodes, d.next);
        std::swap(d.nid, d.nid_next);
        nn = c.nn_next;
        if ((size_t) nn > node_cap) { cuda_ok = false; err = "GPU builder: level table overflow"; break; }
    }
    if (cuda_ok) {
        stats.heap_ranges = c.heaps;
        stats.host_fallback_nodes = c.fallback;
        nodes_out.resize(c.n_top);
        order.resize(n);
        const auto t_k1 = std::chrono::steady_clock::now();
        stats.kernel_seconds = std::chrono::duration<double>(t_k1 - t_kernels).count();
        chk(cudaMemcpyAsync(nodes_out.data(), d.top, (size_t) c.n_top * sizeof(GpuTopNode), cudaMemcpyDeviceToHost, st));
        chk(cudaMemcpyAsync(order.data(), d.idx, (size_t) n * sizeof(uint32_t), cudaMemcpyDeviceToHost, st));
        chk(cudaStreamSynchronize(st));
    }
    if (!cuda_ok) cudaGetLastError();
    stats.seconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - t_begin).count();
    return cuda_ok;
}

}  // namespace rtb
