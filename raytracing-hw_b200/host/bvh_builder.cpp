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

#include <algorithm>
#include <atomic>
#include <chrono>
#include <cstring>
#include <limits>
#include <stdexcept>
#include <omp.h>

namespace rtb {
namespace {

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
};

struct Builder {
    size_t n = 0;
    std::vector<Box> tbox;           // Primitive::aabb() cache (primitive.cpp:63-75)
    std::vector<float> tcenter;      // n x 3
    std::vector<uint32_t> order;
    std::vector<TNode> pool;
    std::atomic<uint32_t> pool_next{0};
    size_t task_cutoff = 1 << 14;

    uint32_t alloc_pair() { return pool_next.fetch_add(2); }

    // buildHelperSAH (bvh.cpp:5-112): returns the partition offset; writes best_dim.
    size_t split(const Box &node_box, size_t first, size_t count, int32_t &best_dim, Scratch &sc) {
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
        for (size_t i = 0; i < count; ++i) {
            const uint32_t t = order[first + i];
            sc.keys[i] = {tcenter[(size_t) t * 3 + dim], t};
        }
        std::sort(sc.keys.begin(), sc.keys.end(), [](const KeyIdx &a, const KeyIdx &b) { return a.key < b.key; });
        for (size_t i = 0; i < count; ++i) order[first + i] = sc.keys[i].idx;

        sc.left_scan.resize(count);
        sc.right_scan.resize(count);
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
        (void) result;
        const int bd = best_dim;
        const float bs = best_split;
        const float *ctr = tcenter.data();
        uint32_t *b = order.data() + first;
        uint32_t *e = b + count;
        uint32_t *mid = std::partition(b, e, [ctr, bd, bs](uint32_t t) { return ctr[(size_t) t * 3 + bd] < bs; });
        return (size_t) (mid - b);
    }

    // BVH::buildNode (bvh.cpp:114-164) for pool node `id`, recursing into the children.
    void build(uint32_t id, size_t first, size_t count, Scratch *sc_in) {
        Scratch local;
        Scratch &sc = sc_in ? *sc_in : local;
        TNode &nd = pool[id];
        nd.box.reset();
        for (size_t i = 0; i < count; ++i) nd.box.grow(tbox[order[first + i]]);
        if (count <= 4) {
            nd.first = (uint32_t) first;
            nd.count = (uint32_t) count;
            return;
        }
        const size_t res = split(nd.box, first, count, nd.split_dim, sc);
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
};

}  // namespace

void HostTriangles::permute(const std::vector<uint32_t> &order) {
    const size_t n = order.size();
    auto apply = [&](auto &vec, size_t stride) {
        using T = typename std::remove_reference<decltype(vec)>::type;
        if (vec.empty()) return;   // (an attribute array that has not been assembled)
        T out(n * stride);
#pragma omp parallel for schedule(static)
        for (size_t i = 0; i < n; ++i)
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

    Builder B;
    B.n = n;
    B.tbox.resize(n);
    B.tcenter.resize(n * 3);
    B.order.resize(n);
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
    B.pool.resize(2 * n + 1);
    B.pool_next = 1;   // slot 0 = root; pairs are allocated from 1

    if (threads <= 0) threads = omp_get_max_threads();
    if (threads <= 1 || n < B.task_cutoff) {
        B.task_cutoff = std::numeric_limits<size_t>::max();
        Scratch sc;
        B.build(0, 0, n, &sc);
    } else {
#pragma omp parallel num_threads(threads)
#pragma omp single
        B.build(0, 0, n, nullptr);
    }

    // Reference numbering: nodes = [root] + desc(root); desc(X) = [L, R] + desc(R) + desc(L).
    const size_t total = 1 + (size_t) B.pool[0].desc;
    out.aabb.resize(total * 6);
    out.links.resize(total * 5);
    struct Item { uint32_t pool_id; uint32_t ref_id; uint32_t desc_base; };
    std::vector<Item> stack;
    stack.push_back({0, 0, 1});
    while (!stack.empty()) {
        const Item it = stack.back();
        stack.pop_back();
        const TNode &t = B.pool[it.pool_id];
        float *a = &out.aabb[(size_t) it.ref_id * 6];
        for (int k = 0; k < 3; ++k) { a[k] = t.box.mn[k]; a[3 + k] = t.box.mx[k]; }
        int32_t *l = &out.links[(size_t) it.ref_id * 5];
        l[2] = t.split_dim;
        if (t.left < 0) {
            l[0] = -1; l[1] = -1;
            l[3] = (int32_t) t.first;
            l[4] = (int32_t) t.count;
        } else {
            const uint32_t L = it.desc_base, R = it.desc_base + 1;
            l[0] = (int32_t) L; l[1] = (int32_t) R;
            l[3] = -1; l[4] = 0;
            const uint32_t r_base = it.desc_base + 2;
            const uint32_t l_base = r_base + B.pool[t.right].desc;
            stack.push_back({(uint32_t) t.left, L, l_base});
            stack.push_back({(uint32_t) t.right, R, r_base});
        }
    }
    tris.permute(B.order);
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
        lights = tris;
        lights.permute(pick);
        build_bvh(lights, light_bvh);
    }
    build_seconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
}

}  // namespace rtb
