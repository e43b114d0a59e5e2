// The per-thread steps of the GPU BVH builder (csrc/rtb_build.cu), written so that the same source runs on the host:
// tests/devsim drives them with plain loops and checks the result against libstdc++'s std::sort itself.
//
// What is being reproduced.  BVH::buildBVH (src/core/bvh.cpp:5-175) sorts the triangle range of every node with
// std::sort on the centroid of the node's widest axis (bvh.cpp:47-50), sweeps the SAH over the sorted range and cuts
// it; the triangle AND node layout is therefore whatever libstdc++'s introsort leaves behind, ties included.
// std::sort(first, last, comp)  =  __introsort_loop(first, last, 2 * lg(n));  __final_insertion_sort(first, last)
// (bits/stl_algo.h).  The loop, on a range of more than 16 elements with depth left:
//      __move_median_to_first(first, first + 1, first + n / 2, last - 1)        // the pivot goes to *first
//      cut = __unguarded_partition(first + 1, last, *first)
//      recurse on [cut, last), carry on with [first, cut)                        // both with depth - 1
// and __unguarded_partition is
//      loop { while (*lo < pivot) ++lo;  --hi;  while (pivot < *hi) --hi;  if (!(lo < hi)) return lo;  swap(*lo, *hi);  ++lo; }
// Neither cursor ever looks at an element the other one has written, so WHERE they stop is a function of the range as
// it was before the first swap:  the left cursor stops at the elements with !(x < pivot), in ascending position
// l_1 < l_2 < ..., the right cursor at those with !(pivot < x), in descending position r_1 > r_2 > ...;  swap k
// exchanges l_k and r_k, and swaps happen for k = 1..S, S = the number of k with l_k < r_k (a prefix: l grows, r falls).
// The cursors meet at  cut = min(l_{S+1}, r_S)  (missing terms = +infinity).  That is two prefix sums, a scatter and an
// exchange per range -- every range of a round in the same passes over the array.  The two parts of a cut never meet
// again, so the recursion is replayed breadth first: one round = one partition step of every range that still has more
// than 16 elements.  What the loop leaves behind are blocks of <= 16 elements with !(y < x) for x in an earlier and y
// in a later block, on which __final_insertion_sort is a STABLE insertion sort of every block (it never moves an
// element out of its block).  Ranges whose depth runs out before they are small are heap-sorted by the reference
// (__partial_sort(first, last, last)); partial_sort_exact below replays libstdc++'s heap primitives move for move.
#pragma once
#include <cstdint>

#ifndef RTB_HD
#if defined(__CUDACC__)
#define RTB_HD __host__ __device__ __forceinline__
#else
#define RTB_HD inline
#endif
#endif

namespace rtb {
namespace gbuild {

constexpr uint32_t kNoRange = 0xFFFFFFFFu;
constexpr uint32_t kInf = 0xFFFFFFFFu;
constexpr int kSmall = 16;   // _S_threshold of bits/stl_algo.h

// std::__lg: floor(log2(n)), n >= 1
RTB_HD int lg_floor(uint32_t n) {
    int l = 0;
    while (n >>= 1) ++l;
    return l;
}

// state of an introsort range
enum : int32_t { kRangeActive = 0, kRangeSmall = 1, kRangeHeap = 2, kRangeDead = 3 };

struct SortRange {
    uint32_t lo, hi;      // [lo, hi)
    int32_t depth;        // depth_limit left
    int32_t state;
};
// per-round results of an active range
struct SortRound {
    float pivot;
    uint32_t cut;
    uint32_t right_id;    // id of the new range [cut, hi)
    uint32_t pad;
};

RTB_HD int32_t range_state(uint32_t lo, uint32_t hi, int32_t depth) {
    if (hi - lo <= (uint32_t) kSmall) return kRangeSmall;
    return depth == 0 ? kRangeHeap : kRangeActive;
}

RTB_HD void swap_at(float *key, uint32_t *idx, uint32_t a, uint32_t b) {
    const float k = key[a];
    key[a] = key[b];
    key[b] = k;
    const uint32_t i = idx[a];
    idx[a] = idx[b];
    idx[b] = i;
}

// __move_median_to_first(result = lo, a = lo + 1, b = lo + n / 2, c = hi - 1) with comp = (key < key)
RTB_HD void median_to_first(float *key, uint32_t *idx, uint32_t lo, uint32_t hi) {
    const uint32_t a = lo + 1, b = lo + (hi - lo) / 2, c = hi - 1;
    const float ka = key[a], kb = key[b], kc = key[c];
    uint32_t pick;
    if (ka < kb) {
        if (kb < kc) pick = b;
        else if (ka < kc) pick = c;
        else pick = a;
    } else if (ka < kc) pick = a;
    else if (kb < kc) pick = c;
    else pick = b;
    swap_at(key, idx, lo, pick);
}

// element i of an active range [lo, hi) with pivot p (lo < i < hi): where the cursors stop
RTB_HD bool stops_left(float k, float pivot) { return !(k < pivot); }
RTB_HD bool stops_right(float k, float pivot) { return !(pivot < k); }

// __insertion_sort(first, last) with comp = (key < key): stable
RTB_HD void insertion_sort_block(float *key, uint32_t *idx, uint32_t lo, uint32_t hi) {
    for (uint32_t i = lo + 1; i < hi; ++i) {
        const float k = key[i];
        const uint32_t v = idx[i];
        uint32_t j = i;
        while (j > lo && k < key[j - 1]) {
            key[j] = key[j - 1];
            idx[j] = idx[j - 1];
            --j;
        }
        key[j] = k;
        idx[j] = v;
    }
}

// ---- libstdc++'s heap primitives on (key, idx) pairs (bits/stl_heap.h), comp = (key < key)
RTB_HD void push_heap_exact(float *key, uint32_t *idx, uint32_t first, int64_t hole, int64_t top, float vk, uint32_t vi) {
    int64_t parent = (hole - 1) / 2;
    while (hole > top && key[first + parent] < vk) {
        key[first + hole] = key[first + parent];
        idx[first + hole] = idx[first + parent];
        hole = parent;
        parent = (hole - 1) / 2;
    }
    key[first + hole] = vk;
    idx[first + hole] = vi;
}
RTB_HD void adjust_heap_exact(float *key, uint32_t *idx, uint32_t first, int64_t hole, int64_t len, float vk, uint32_t vi) {
    const int64_t top = hole;
    int64_t child = hole;
    while (child < (len - 1) / 2) {
        child = 2 * (child + 1);
        if (key[first + child] < key[first + child - 1]) --child;
        key[first + hole] = key[first + child];
        idx[first + hole] = idx[first + child];
        hole = child;
    }
    if ((len & 1) == 0 && child == (len - 2) / 2) {
        child = 2 * (child + 1);
        key[first + hole] = key[first + child - 1];
        idx[first + hole] = idx[first + child - 1];
        hole = child - 1;
    }
    push_heap_exact(key, idx, first, hole, top, vk, vi);
}
// std::__partial_sort(first, last, last): __heap_select(first, last, last) = __make_heap(first, last) (the loop over
// [middle, last) is empty), then __sort_heap(first, last)
RTB_HD void partial_sort_exact(float *key, uint32_t *idx, uint32_t lo, uint32_t hi) {
    const int64_t len = (int64_t) hi - (int64_t) lo;
    if (len < 2) return;
    for (int64_t parent = (len - 2) / 2;; --parent) {   // __make_heap
        const float vk = key[lo + parent];
        const uint32_t vi = idx[lo + parent];
        adjust_heap_exact(key, idx, lo, parent, len, vk, vi);
        if (parent == 0) break;
    }
    for (int64_t last = len; last > 1;) {               // __sort_heap: __pop_heap(first, last - 1, last - 1)
        --last;
        const float vk = key[lo + last];
        const uint32_t vi = idx[lo + last];
        key[lo + last] = key[lo];
        idx[lo + last] = idx[lo];
        adjust_heap_exact(key, idx, lo, 0, last, vk, vi);
    }
}

// ---- AABB arithmetic in the reference's order (primitive.h:22-51, vector.h:326-328)
struct Box6 {
    float mn[3], mx[3];
};
// AABB::grow(AABB): min = std::min(min, b.min) -> (b < a) ? b : a;  max = std::max(max, b.max) -> (a < b) ? b : a
// (`a` is the running box: of two equal bounds the EARLIER one stays, which decides the sign of a zero)
RTB_HD Box6 box_grow(const Box6 &a, const Box6 &b) {
    Box6 r;
    for (int k = 0; k < 3; ++k) {
        r.mn[k] = (b.mn[k] < a.mn[k]) ? b.mn[k] : a.mn[k];
        r.mx[k] = (a.mx[k] < b.mx[k]) ? b.mx[k] : a.mx[k];
    }
    return r;
}
RTB_HD float box_size(const Box6 &b, int k) { return b.mx[k] + (-b.mn[k]); }
RTB_HD float box_area(const Box6 &b) {
    const float x = box_size(b, 0), y = box_size(b, 1), z = box_size(b, 2);
    return 2.f * (x * y + x * z + y * z);
}
// widest axis, first of equals (bvh.cpp:26-34)
RTB_HD int box_widest(const Box6 &b) {
    int dim = 0;
    float max_size = box_size(b, 0);
    for (int i = 0; i < 3; ++i)
        if (max_size < box_size(b, i)) { max_size = box_size(b, i); dim = i; }
    return dim;
}
// metric of cutting a sorted range of `count` after its first `left` triangles (bvh.cpp:80-84)
RTB_HD float sah_metric(float left_area, float right_area, uint32_t left, uint32_t count) {
    return left_area * (float) left + right_area * (float) (count - left);
}
// a non-negative float's bits order like the float
RTB_HD unsigned long long pack_metric(float m, uint32_t i) {
    union { float f; uint32_t u; } c;
    c.f = m;
    return ((unsigned long long) c.u << 32) | (unsigned long long) i;
}

}  // namespace gbuild
}  // namespace rtb
