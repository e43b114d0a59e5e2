// Host BVH builder whose node AND triangle layout equals the reference's (BVH::buildBVH, src/core/bvh.cpp:5-175).
//
// The reference sorts 208-byte Primitive structs with std::sort on the widest axis at every node, does a full
// SAH sweep (BIN_SIZE 1), partitions with std::partition, and numbers nodes in the order a LIFO work list
// processes them (right child first).  Layout therefore depends on libstdc++'s unstable sort/partition move
// sequence.  We run the SAME libstdc++ algorithms with the SAME comparator outcomes on
// (key, index) pairs / an index array -- the permutation produced by std::sort and std::partition is a function of
// the comparison results only -- and reproduce the numbering analytically so that sub-trees can be built in
// parallel:  nodes = [root] + desc(root),  desc(X) = [L, R] + desc(R) + desc(L)   (bvh.cpp:156-161 pushes L then R,
// bvh.cpp:115-116 pops the back).
//
// All float arithmetic restates the reference's op order: AABB::grow via std::min/std::max argument order
// (primitive.h:22-30), size = max + (-min) (vector.h:326-328), center = min + size*0.5f (primitive.h:43-45),
// surface_area = 2*((xy + xz) + yz) (primitive.h:48-51).  This TU must be compiled WITHOUT -ffast-math and
// without FMA contraction (-ffp-contract=off), like the reference's x86-64 baseline build.
#include "host_scene.h"
#include "../csrc/rtb_build.h"

#include <algorithm>
#include <atomic>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <new>
#include <cstring>
#include <limits>
#include <stdexcept>
#include <omp.h>
#include <sys/mman.h>

namespace rtb {
namespace {
GpuBuildFn g_gpu_build = &gpu_build_top;
}
void set_gpu_builder_for_tests(GpuBuildFn fn) { g_gpu_build = fn ? fn : &gpu_build_top; }
namespace {

// Large work arrays are first touched right after they are allocated; with 4 KB pages that is ~100 k page faults per array
// at 10 M triangles (0.1 - 0.5 s each, and they serialise across threads).  Ask for transparent huge pages on the part of
// the block that is 2 MB aligned (a no-op where the kernel does not offer them).
inline void advise_huge(void *p, size_t bytes) {
    const uintptr_t a = ((uintptr_t) p + ((uintptr_t) 2 << 20) - 1) & ~(((uintptr_t) 2 << 20) - 1);
    const uintptr_t e = ((uintptr_t) p + bytes) & ~(((uintptr_t) 2 << 20) - 1);
    if (e > a) madvise((void *) a, (size_t) (e - a), MADV_HUGEPAGE);
}
template <class V>
void resize_huge(V &v, size_t count) {
    v.reserve(count);
    if (count * sizeof(v[0]) >= ((size_t) 8 << 20)) advise_huge(v.data(), count * sizeof(v[0]));
    v.resize(count);
}

struct Box {
    float mn[3], mx[3];
    void reset() {
        const float m = std::numeric_limits<float>::max();
        for (int k = 0; k < 3; ++k) { mn[k] = m; mx[k] = -m; }
    }
    // AABB::grow(vector3f) / grow(AABB): min = ::min(min, p) -> std::min(min.x, p.x) -> (p < min) ? p : min
    void grow_point(const float *p) {
        for (int k = 0; k < 3; ++k) {
            mn[k] = (p[k] < mn[k]) ? p[k] : mn[k];
            mx[k] = (mx[k] < p[k]) ? p[k] : mx[k];
        }
    }
    void grow(const Box &b) {
        for (int k = 0; k < 3; ++k) {
            mn[k] = (b.mn[k] < mn[k]) ? b.mn[k] : mn[k];
            mx[k] = (mx[k] < b.mx[k]) ? b.mx[k] : mx[k];
        }
    }
    float size(int k) const { return mx[k] + (-mn[k]); }
    float center(int k) const { return mn[k] + size(k) * 0.5f; }
    float surface_area() const {
        const float x = size(0), y = size(1), z = size(2);
        return 2.f * (x * y + x * z + y * z);
    }
};

struct TNode {
    Box box;
    int32_t left = -1, right = -1;   // pool indices
    int32_t split_dim = -1;
    uint32_t first = 0, count = 0;   // leaf range (count == 0 for internal nodes)
    uint32_t desc = 0;               // |desc(X)|
};

struct KeyIdx {
    float key;
    uint32_t idx;
};

struct Scratch {
    std::vector<KeyIdx> keys;
    std::vector<float> left_scan, right_scan;
    std::vector<Box> boxes;   // wide nodes: the triangles' boxes in sorted order (one random gather instead of three)
};


// ---- std::sort, exactly, on several threads.  libstdc++'s std::sort(first, last, comp) is
//   __introsort_loop(first, last, 2 * lg(n), comp);  __final_insertion_sort(first, last, comp);      (bits/stl_algo.h)
// The loop cuts [first, last) at __unguarded_partition_pivot and recurses into the RIGHT part, iterating on the left; the
// two parts never touch each other again, so they can be replayed as independent tasks -- the sequence of comparisons
// and moves inside each part is the sequential one.  What the loop leaves behind is a sequence of blocks of <= 16
// elements (or heap-sorted ranges where the depth limit ran out) with  !(y < x)  for x in an earlier, y in a later block;
// on that the final insertion sort never moves an element out of its block and is a stable insertion sort of every
// block, which is what __insertion_sort of the block does on the spot.  Same permutation as std::sort for every input
// (ties included): tests/test_layout.py builds every scene with the sequential and the task-parallel sort.
// RTB_BUILD_SORT_DEPTH=<d>: introsort's depth limit (2 * lg(n)) replaced by d in every sort of the build, host and GPU
// alike -- lets the tests reach the heap-sorted ranges that real inputs almost never produce
// (read once per build_bvh call)
int g_sort_depth = -1;
int sort_depth_override() { return g_sort_depth; }
inline auto key_comp() { return __gnu_cxx::__ops::__iter_comp_iter([](const KeyIdx &a, const KeyIdx &b) { return a.key < b.key; }); }

void exact_sort_loop(KeyIdx *first, KeyIdx *last, long depth_limit, size_t task_min) {
    auto comp = key_comp();
    while (last - first > 16) {
        if (depth_limit == 0) {
            std::__partial_sort(first, last, last, comp);   // heapsort, as __introsort_loop falls back to
            return;
        }
        --depth_limit;
        KeyIdx *cut = std::__unguarded_partition_pivot(first, last, comp);
        if ((size_t) (last - cut) >= task_min) {
#pragma omp task firstprivate(cut, last, depth_limit, task_min)
            exact_sort_loop(cut, last, depth_limit, task_min);
        } else {
            exact_sort_loop(cut, last, depth_limit, task_min);
        }
        last = cut;
    }
    std::__insertion_sort(first, last, comp);
}
// must be called from inside a parallel region (tasks); returns when the range is sorted
void exact_sort_parallel(KeyIdx *first, KeyIdx *last, size_t task_min) {
    if (first == last) return;
    const int ov = sort_depth_override();
#pragma omp taskgroup
    exact_sort_loop(first, last, ov >= 0 ? (long) ov : (long) std::__lg(last - first) * 2, task_min);
}
// std::sort, or the same loops with the depth limit of sort_depth_override()
void exact_sort_serial(KeyIdx *first, KeyIdx *last) {
    const int ov = sort_depth_override();
    if (ov < 0) {
        std::sort(first, last, [](const KeyIdx &a, const KeyIdx &b) { return a.key < b.key; });
    } else if (first != last) {
        auto comp = key_comp();
        std::__introsort_loop(first, last, (long) ov, comp);
        std::__final_insertion_sort(first, last, comp);
    }
}

// i-th of `parts` equal chunks of [0, n)
inline size_t chunk_at(size_t n, size_t parts, size_t i) { return n / parts * i + std::min(i, n % parts); }

struct Builder {
    size_t n = 0;
    std::vector<Box> tbox;           // Primitive::aabb() cache (primitive.cpp:63-75)
    std::vector<float> tcenter;      // n x 3
    std::vector<uint32_t> order;
    // node pool: 2 n + 1 slots of address space, touched only where nodes are made (value-initialising 2 n + 1 records up
    // front cost 0.5 s of page faults at 10 M triangles)
    TNode *pool = nullptr;
    std::atomic<uint32_t> pool_next{0};
    ~Builder() { std::free(pool); }
    size_t task_cutoff = 1 << 14;

    uint32_t alloc_pair() {
        const uint32_t p = pool_next.fetch_add(2);
        new (&pool[p]) TNode();
        new (&pool[p + 1]) TNode();
        return p;
    }

    size_t wide_cutoff = (size_t) 1 << 17;   // nodes at least this large are worked on by all threads (taskloops)
    size_t sort_task_min = (size_t) 1 << 15; // a partition's right part of at least this size becomes a task
    int wide_parts = 64;

    // union of the triangles' boxes in array order (bvh.cpp:130-132); chunked, combined in order: AABB::grow keeps the
    // EARLIER of two equal bounds (strict compares), so the in-order combination of in-order chunks gives the same bits
    Box range_box(size_t first, size_t count, bool wide) {
        Box out;
        out.reset();
        if (!wide) {
            for (size_t i = 0; i < count; ++i) out.grow(tbox[order[first + i]]);
            return out;
        }
        const size_t P = (size_t) wide_parts;
        std::vector<Box> part(P);
#pragma omp taskloop default(shared) grainsize(1)
        for (size_t c = 0; c < P; ++c) {
            Box b;
            b.reset();
            for (size_t i = chunk_at(count, P, c), e = chunk_at(count, P, c + 1); i < e; ++i) b.grow(tbox[order[first + i]]);
            part[c] = b;
        }
        for (size_t c = 0; c < P; ++c)
            if (chunk_at(count, P, c) < chunk_at(count, P, c + 1)) out.grow(part[c]);
        return out;
    }

    // buildHelperSAH (bvh.cpp:5-112): returns the partition offset; writes best_dim.
    size_t split(const Box &node_box, size_t first, size_t count, int32_t &best_dim, Scratch &sc, bool wide) {
        float best_sah = node_box.surface_area() * (float) count;
        // `float best_split = center[2] - 1e-7;` : float minus double literal, rounded back to float (bvh.cpp:20)
        float best_split = (float) ((double) tcenter[(size_t) order[first] * 3 + 2] - 1e-7);
        size_t result = 0;
        best_dim = 2;

        int dim = 0;
        {
            float max_size = node_box.size(0);
            for (int i = 0; i < 3; ++i)
                if (max_size < node_box.size(i)) { max_size = node_box.size(i); dim = i; }
        }

        sc.keys.resize(count);
        sc.left_scan.resize(count);
        sc.right_scan.resize(count);
        if (!wide) {
            for (size_t i = 0; i < count; ++i) {
                const uint32_t t = order[first + i];
                sc.keys[i] = {tcenter[(size_t) t * 3 + dim], t};
            }
            exact_sort_serial(sc.keys.data(), sc.keys.data() + count);
            for (size_t i = 0; i < count; ++i) order[first + i] = sc.keys[i].idx;

            Box tmp = tbox[order[first]];
            sc.left_scan[0] = tmp.surface_area();
            for (size_t i = 1; i < count; ++i) {
                tmp.grow(tbox[order[first + i]]);
                sc.left_scan[i] = tmp.surface_area();
            }
            tmp = tbox[order[first + count - 1]];
            sc.right_scan[count - 1] = tmp.surface_area();
            for (size_t i = count - 1; i-- > 0;) {
                tmp.grow(tbox[order[first + i]]);
                sc.right_scan[i] = tmp.surface_area();
            }
            for (size_t i = 0; i + 1 < count; ++i) {
                const float ls = sc.left_scan[i];
                const float rs = sc.right_scan[i + 1];
                const size_t left_count = i + 1;
                const float metric = ls * (float) left_count + rs * (float) (count - left_count);
                if (metric < best_sah) {
                    best_sah = metric;
                    result = left_count;
                    best_dim = dim;
                    best_split = (sc.keys[result - 1].key + sc.keys[result].key) * 0.5f;
                }
            }
        } else {
            // the same, every pass cut into chunks that all threads work on (the sort: exact_sort_parallel)
            const size_t P = (size_t) wide_parts;
            KeyIdx *keys = sc.keys.data();
            uint32_t *ord = order.data() + first;
            const float *ctr = tcenter.data();
#pragma omp taskloop default(shared) grainsize(1)
            for (size_t c = 0; c < P; ++c)
                for (size_t i = chunk_at(count, P, c), e = chunk_at(count, P, c + 1); i < e; ++i) keys[i] = {ctr[(size_t) ord[i] * 3 + dim], ord[i]};
            exact_sort_parallel(keys, keys + count, sort_task_min);
            // prefix / suffix boxes: chunk unions first, then every chunk scans from the union of the chunks before / after it
            std::vector<Box> pre(P), suf(P);
            sc.boxes.resize(count);
            Box *sorted = sc.boxes.data();
#pragma omp taskloop default(shared) grainsize(1)
            for (size_t c = 0; c < P; ++c) {
                const size_t b = chunk_at(count, P, c), e = chunk_at(count, P, c + 1);
                Box fwd, bwd;
                fwd.reset();
                bwd.reset();
                for (size_t i = b; i < e; ++i) { ord[i] = keys[i].idx; sorted[i] = tbox[keys[i].idx]; fwd.grow(sorted[i]); }
                for (size_t i = e; i-- > b;) bwd.grow(sorted[i]);
                pre[c] = fwd;
                suf[c] = bwd;
            }
            std::vector<Box> before(P), after(P);   // union of everything left of chunk c (in order) / right of it (in reverse order)
            {
                Box acc;
                acc.reset();
                for (size_t c = 0; c < P; ++c) { before[c] = acc; if (chunk_at(count, P, c) < chunk_at(count, P, c + 1)) acc.grow(pre[c]); }
                acc.reset();
                for (size_t c = P; c-- > 0;) { after[c] = acc; if (chunk_at(count, P, c) < chunk_at(count, P, c + 1)) acc.grow(suf[c]); }
            }
            float *ls = sc.left_scan.data(), *rs = sc.right_scan.data();
#pragma omp taskloop default(shared) grainsize(1)
            for (size_t c = 0; c < P; ++c) {
                const size_t b = chunk_at(count, P, c), e = chunk_at(count, P, c + 1);
                if (b == e) continue;
                // (the reference starts its running box AS the first box -- `AABB tmp = sah_bins[0]` -- which equals growing
                // an empty box by it: every bound of an empty box loses the strict compare)
                Box tmp = before[c];
                for (size_t i = b; i < e; ++i) { tmp.grow(sorted[i]); ls[i] = tmp.surface_area(); }
                tmp = after[c];
                for (size_t i = e; i-- > b;) { tmp.grow(sorted[i]); rs[i] = tmp.surface_area(); }
            }
            // first index whose metric is strictly below everything before it and below the node's own cost
            struct Best { float v; size_t left; bool any; };
            std::vector<Best> bests(P);
            const float start = best_sah;
#pragma omp taskloop default(shared) grainsize(1)
            for (size_t c = 0; c < P; ++c) {
                Best bb{start, 0, false};
                for (size_t i = chunk_at(count, P, c), e = std::min(chunk_at(count, P, c + 1), count - 1); i < e; ++i) {
                    const size_t left_count = i + 1;
                    const float metric = ls[i] * (float) left_count + rs[i + 1] * (float) (count - left_count);
                    if (metric < bb.v) bb = Best{metric, left_count, true};
                }
                bests[c] = bb;
            }
            for (size_t c = 0; c < P; ++c)
                if (bests[c].any && bests[c].v < best_sah) {
                    best_sah = bests[c].v;
                    result = bests[c].left;
                    best_dim = dim;
                    best_split = (keys[result - 1].key + keys[result].key) * 0.5f;
                }
        }
        (void) result;
        const int bd = best_dim;
        const float bs = best_split;
        const float *ctr = tcenter.data();
        uint32_t *b = order.data() + first;
        uint32_t *e = b + count;
        if (wide && bd == dim) {
            // the range is sorted by the partition key: the predicate is true on a prefix, std::partition's cursors meet
            // at its end without a swap (bits/stl_algo.h, bidirectional __partition) -- find that end directly
            const KeyIdx *keys = sc.keys.data();
            return (size_t) (std::partition_point(keys, keys + count, [bs](const KeyIdx &k) { return k.key < bs; }) - keys);
        }
        uint32_t *mid = std::partition(b, e, [ctr, bd, bs](uint32_t t) { return ctr[(size_t) t * 3 + bd] < bs; });
        return (size_t) (mid - b);
    }

    // BVH::buildNode (bvh.cpp:114-164) for pool node `id`, recursing into the children.
    void build(uint32_t id, size_t first, size_t count, Scratch *sc_in) {
        Scratch local;
        Scratch &sc = sc_in ? *sc_in : local;
        TNode &nd = pool[id];
        const bool wide = count >= wide_cutoff && task_cutoff != std::numeric_limits<size_t>::max();
        nd.box = range_box(first, count, wide);
        if (count <= 4) {
            nd.first = (uint32_t) first;
            nd.count = (uint32_t) count;
            return;
        }
        const size_t res = split(nd.box, first, count, nd.split_dim, sc, wide);
        if (res == 0 || res == count) {
            nd.first = (uint32_t) first;
            nd.count = (uint32_t) count;
            return;
        }
        const uint32_t pair = alloc_pair();
        nd.left = (int32_t) pair;
        nd.right = (int32_t) pair + 1;
        const bool spawn = count >= task_cutoff;
        if (spawn) {
#pragma omp task firstprivate(pair, first, res, count)
            build(pair + 1, first + res, count - res, nullptr);
#pragma omp task firstprivate(pair, first, res)
            build(pair, first, res, nullptr);
#pragma omp taskwait
        } else {
            build(pair + 1, first + res, count - res, &sc);
            build(pair, first, res, &sc);
        }
        TNode &me = pool[id];
        me.desc = 2 + pool[pair].desc + pool[pair + 1].desc;
    }

    // The part of the tree the GPU built (csrc/rtb_build.cu) becomes pool nodes -- with the SAME indices (the root is 0 and
    // the children of a node are a pair in both) -- in one linear pass; the ranges it handed back are built here, in
    // parallel; the subtree sizes then follow from one backward sweep (children have larger indices than their parents).
    void adopt_all(const std::vector<GpuTopNode> &top) {
        const size_t nt = top.size();
        pool_next = (uint32_t) nt;
        std::vector<uint32_t> todo;
#pragma omp parallel
        {
            std::vector<uint32_t> mine;
#pragma omp for schedule(static) nowait
            for (size_t t = 0; t < nt; ++t) {
                const GpuTopNode &g = top[t];
                TNode *nd = new (&pool[t]) TNode();
                if (g.kind == kTopHost) { mine.push_back((uint32_t) t); continue; }
                for (int k = 0; k < 3; ++k) { nd->box.mn[k] = g.box[k]; nd->box.mx[k] = g.box[3 + k]; }
                nd->split_dim = g.split_dim;
                if (g.kind == kTopLeaf) {
                    nd->first = g.first;
                    nd->count = g.count;
                } else {
                    nd->left = g.left;
                    nd->right = g.right;
                }
            }
#pragma omp critical
            todo.insert(todo.end(), mine.begin(), mine.end());
        }
        // Large ranges first (they spawn tasks of their own); then a fixed number of tasks, each taking every T-th range of the
        // size-ordered list.  (A taskloop with a grain size made 1400 tasks for 22 k ranges, and libgomp runs a taskloop of more
        // than 64 x threads tasks serially on the encountering thread: the ranges the GPU builder handed over at 64 - 256
        // triangles were built by ONE thread -- 1.7 s at 10 M triangles where 4096 took 0.6 s.)
        std::sort(todo.begin(), todo.end(), [&](uint32_t x, uint32_t y) { return top[x].count != top[y].count ? top[x].count > top[y].count : x < y; });
#pragma omp parallel
#pragma omp single
        {
            const size_t T = std::min<size_t>(todo.size(), (size_t) omp_get_num_threads() * 16);
#pragma omp taskloop default(shared) num_tasks(T)
            for (size_t k = 0; k < T; ++k)
                for (size_t i = k; i < todo.size(); i += T) {
                    const GpuTopNode &g = top[todo[i]];
                    build(todo[i], g.first, g.count, nullptr);
                }
        }
        for (size_t t = nt; t-- > 0;) {
            TNode &nd = pool[t];
            if (top[t].kind == kTopInternal) nd.desc = 2 + pool[nd.left].desc + pool[nd.right].desc;
        }
    }
};

}  // namespace

// (called from inside a parallel region: the loops are taskloops)
void HostTriangles::permute(const std::vector<uint32_t> &order) {
    const size_t n = order.size();
    auto apply = [&](auto &vec, size_t stride) {
        using T = typename std::remove_reference<decltype(vec)>::type;
        if (vec.empty()) return;   // (an attribute array that has not been assembled)
        T out;
        resize_huge(out, n * stride);
        const size_t parts = std::max<size_t>(1, std::min<size_t>((n + 65535) / 65536, (size_t) omp_get_num_threads() * 32));   // (libgomp serialises a taskloop of > 64 x threads tasks)
#pragma omp taskloop default(shared) num_tasks(parts)
        for (size_t c = 0; c < parts; ++c)
            for (size_t i = chunk_at(n, parts, c), e = chunk_at(n, parts, c + 1); i < e; ++i)
                std::memcpy(&out[i * stride], &vec[(size_t) order[i] * stride], stride * sizeof(out[0]));
        vec.swap(out);
    };
    apply(pos, 9);
    apply(nrm, 9);
    apply(tan, 12);
    apply(uv, 6);
    apply(mesh, 1);
    apply(source, 1);
}

void build_bvh(HostTriangles &tris, HostBvh &out, int threads) {
    const size_t n = tris.size();
    out.aabb.clear();
    out.links.clear();
    if (n == 0) throw std::runtime_error("No primitives for node 0");   // bvh.cpp:122-123
    if (n >= (size_t) 1 << 30) throw std::runtime_error("too many triangles");

    {
        const char *e = std::getenv("RTB_BUILD_SORT_DEPTH");
        g_sort_depth = e && *e ? std::max(0, std::atoi(e)) : -1;
    }
    const bool verbose = std::getenv("RTB_BUILD_VERBOSE") != nullptr;
    auto t_mark = std::chrono::steady_clock::now();
    auto lap = [&](const char *what) {
        if (!verbose) return;
        const auto now = std::chrono::steady_clock::now();
        std::fprintf(stderr, "[build_bvh] %-28s %.3f s\n", what, std::chrono::duration<double>(now - t_mark).count());
        t_mark = now;
    };
    Builder B;
    B.n = n;
    resize_huge(B.tbox, n);
    resize_huge(B.tcenter, n * 3);
    resize_huge(B.order, n);
#pragma omp parallel for schedule(static)
    for (size_t i = 0; i < n; ++i) {
        const float *p = &tris.pos[i * 9];
        float pts[3][3];
        for (int k = 0; k < 3; ++k) {
            pts[0][k] = p[k];
            pts[1][k] = p[k] + p[3 + k];     // points[1] = position[0] + position[1]  (primitive.cpp:67-68)
            pts[2][k] = p[k] + p[6 + k];
        }
        Box b;
        b.reset();
        for (auto &q : pts) b.grow_point(q);
        B.tbox[i] = b;
        for (int k = 0; k < 3; ++k) B.tcenter[i * 3 + k] = b.center(k);
        B.order[i] = (uint32_t) i;
    }
    B.pool = static_cast<TNode *>(std::malloc((2 * n + 1) * sizeof(TNode)));
    if (!B.pool) throw std::runtime_error("out of memory for the BVH node pool");
    if (n >= ((size_t) 1 << 18)) advise_huge(B.pool, (2 * n + 1) * sizeof(TNode));
    new (&B.pool[0]) TNode();
    B.pool_next = 1;   // slot 0 = root; pairs are allocated from 1

    lap("triangle boxes / centres");
    if (threads <= 0) threads = omp_get_max_threads();
    // RTB_BUILD_WIDE=<n>: nodes of >= n triangles take the all-threads path (tests force it onto small scenes)
    if (const char *e = std::getenv("RTB_BUILD_WIDE")) {
        const long v = std::atol(e);
        if (v > 0) {
            B.wide_cutoff = (size_t) v;
            B.sort_task_min = std::max<size_t>(17, (size_t) v / 8);
            B.task_cutoff = std::min(B.task_cutoff, std::max<size_t>(64, (size_t) v / 2));
            B.wide_parts = 7;
        }
    }
    // RTB_BUILD_GPU: 0 never, 1 whenever a device is there, unset: scenes of >= 500 k triangles.  RTB_BUILD_GPU_BELOW: nodes with
    // fewer triangles than this are built on the host (>= 5; 5, the default = the whole tree on the GPU).
    bool on_gpu = false;
    {
        const char *e = std::getenv("RTB_BUILD_GPU");
        const int mode = e && *e ? std::atoi(e) : -1;
        size_t below = 5;   // the whole tree: measured fastest (profiles/r02g_build_time.jsonl)
        if (const char *b = std::getenv("RTB_BUILD_GPU_BELOW")) below = (size_t) std::max(5L, std::atol(b));
        if (mode != 0 && (mode > 0 || n >= (size_t) 500000) && n >= below && threads != 1) {
            std::vector<GpuTopNode> top;
            std::vector<uint32_t> ord;
            GpuBuildStats gs;
            std::string err;
            if (g_gpu_build(reinterpret_cast<const float *>(B.tbox.data()), B.tcenter.data(), n, below, -1, sort_depth_override(), top, ord, gs, err)) {
                B.order.swap(ord);
                out.gpu = gs;
                out.built_on_gpu = true;
                lap("GPU levels (incl. copies)");
                B.adopt_all(top);
                on_gpu = true;
            } else if (mode > 0 && std::getenv("RTB_BUILD_GPU_STRICT")) {
                throw std::runtime_error("GPU BVH build failed: " + err);
            } else if (err != "no CUDA device") {
                // (same layout either way, but seconds slower on large scenes: say so once per build)
                std::fprintf(stderr, "rtb: the GPU BVH build was discarded (%s); building on the host\n", err.c_str());
            }
        }
    }
    if (on_gpu) {
    } else if (threads <= 1 || n < B.task_cutoff) {
        B.task_cutoff = std::numeric_limits<size_t>::max();
        Scratch sc;
        B.build(0, 0, n, &sc);
    } else {
#pragma omp parallel num_threads(threads)
#pragma omp single
        B.build(0, 0, n, nullptr);
    }

    lap("host nodes");
    // Reference numbering: nodes = [root] + desc(root); desc(X) = [L, R] + desc(R) + desc(L).
    const size_t total = 1 + (size_t) B.pool[0].desc;
    resize_huge(out.aabb, total * 6);
    resize_huge(out.links, total * 5);
    // (the number of every node follows from the sizes of the subtrees alone, so the subtrees are numbered in parallel)
    struct Numberer {
        const Builder &B;
        HostBvh &out;
        void run(uint32_t pool_id, uint32_t ref_id, uint32_t desc_base) const {
            while (true) {
                const TNode &t = B.pool[pool_id];
                float *a = &out.aabb[(size_t) ref_id * 6];
                for (int k = 0; k < 3; ++k) { a[k] = t.box.mn[k]; a[3 + k] = t.box.mx[k]; }
                int32_t *l = &out.links[(size_t) ref_id * 5];
                l[2] = t.split_dim;
                if (t.left < 0) {
                    l[0] = -1; l[1] = -1;
                    l[3] = (int32_t) t.first;
                    l[4] = (int32_t) t.count;
                    return;
                }
                const uint32_t L = desc_base, R = desc_base + 1;
                l[0] = (int32_t) L; l[1] = (int32_t) R;
                l[3] = -1; l[4] = 0;
                const uint32_t r_base = desc_base + 2;
                const uint32_t l_base = r_base + B.pool[t.right].desc;
                const uint32_t right = (uint32_t) t.right, left = (uint32_t) t.left;
                if (t.desc >= 8192) {
#pragma omp task firstprivate(right, R, r_base)
                    run(right, R, r_base);
                } else {
                    run(right, R, r_base);
                }
                pool_id = left;
                ref_id = L;
                desc_base = l_base;
            }
        }
    };
    {
        const Numberer nb{B, out};
#pragma omp parallel num_threads(threads)
#pragma omp single
        {
#pragma omp taskgroup
            {
#pragma omp task default(shared)
                nb.run(0, 0, 1);
                tris.permute(B.order);   // independent of the numbering: both run on the same team
            }
        }
    }
    lap("numbering + triangle permutation");
}

void HostScene::finalize() {
    auto t0 = std::chrono::steady_clock::now();
    build_bvh(tris, bvh);
    // rng::ManyLightsDistribution (random.cpp:156-167): copy the emissive triangles in post-build order, then
    // build a second BVH over the copy (which reorders the copy).  emissive <=> material.emission != (0,0,0).
    lights = HostTriangles();
    light_bvh = HostBvh();
    std::vector<uint32_t> pick;
    for (size_t i = 0; i < tris.size(); ++i) {
        const float *m = &mesh_material[(size_t) tris.mesh[i] * 16];
        if (!(m[5] == 0.f && m[6] == 0.f && m[7] == 0.f)) pick.push_back((uint32_t) i);
    }
    if (!pick.empty()) {
        auto take = [&](const auto &src, auto &dst, size_t stride) {   // lights[i] = tris[pick[i]]
            if (src.empty()) return;
            dst.resize(pick.size() * stride);
            for (size_t i = 0; i < pick.size(); ++i) std::memcpy(&dst[i * stride], &src[(size_t) pick[i] * stride], stride * sizeof(dst[0]));
        };
        take(tris.pos, lights.pos, 9); take(tris.nrm, lights.nrm, 9); take(tris.tan, lights.tan, 12); take(tris.uv, lights.uv, 6);
        take(tris.mesh, lights.mesh, 1); take(tris.source, lights.source, 1);
        build_bvh(lights, light_bvh);
    }
    build_seconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
}

}  // namespace rtb
