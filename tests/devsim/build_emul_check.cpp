// Test tooling: the GPU BVH builder ITSELF (csrc/rtb_build.cu: kernels and level loop, unchanged) compiled for the host -- every
// kernel launch runs its threads one after the other over host memory (emul/cuda_emulation.h; the Makefile rewrites the <<<>>>
// launches) -- and compared, node for node and triangle for triangle, with the host builder behind rtb_build_bvh (librtb.so,
// RTB_BUILD_GPU=0), which tests/test_layout.py pins against the compiled reference.  Covers what the GPU tests cannot afford
// or do not reach: degenerate geometry (coincident triangles, flat axis-aligned grids with zeros of both signs), nodes no cut
// improves (std::partition on z replayed in the node's thread; handed back above 256 triangles), tiny scenes, forced depth limits.
#define RTB_BUILD_EMULATE 1
#define gpu_build_top gpu_build_top_emulated   // (librtb.so has the real one; this is rtb::gpu_build_top_emulated)
#include "build_emul_gen.inc"
#undef gpu_build_top
namespace rtb_emul = rtb;

#include <cstdio>
#include <limits>
#include <random>

extern "C" int rtb_build_bvh(const float *tri_pos9, size_t n, uint32_t *order, float *node_aabb, int32_t *node_links, size_t node_capacity,
                             size_t *n_nodes, double out8[8]);
extern "C" const char *rtb_last_error(void);

using rtb_emul::GpuTopNode;

// Primitive::aabb() / center() as host/bvh_builder.cpp computes them (primitive.cpp:63-75, primitive.h:43-45)
static void boxes_of(const std::vector<float> &pos, std::vector<float> &tbox, std::vector<float> &ctr) {
    const size_t n = pos.size() / 9;
    tbox.resize(n * 6);
    ctr.resize(n * 3);
    for (size_t i = 0; i < n; ++i) {
        const float *p = &pos[i * 9];
        float pts[3][3];
        for (int k = 0; k < 3; ++k) { pts[0][k] = p[k]; pts[1][k] = p[k] + p[3 + k]; pts[2][k] = p[k] + p[6 + k]; }
        float mn[3], mx[3];
        const float m = std::numeric_limits<float>::max();
        for (int k = 0; k < 3; ++k) { mn[k] = m; mx[k] = -m; }
        for (auto &q : pts)
            for (int k = 0; k < 3; ++k) { mn[k] = (q[k] < mn[k]) ? q[k] : mn[k]; mx[k] = (mx[k] < q[k]) ? q[k] : mx[k]; }
        for (int k = 0; k < 3; ++k) {
            tbox[i * 6 + k] = mn[k];
            tbox[i * 6 + 3 + k] = mx[k];
            const float size = mx[k] + (-mn[k]);
            ctr[i * 3 + k] = mn[k] + size * 0.5f;
        }
    }
}

// the reference's numbering of the emulated builder's nodes: nodes = [root] + desc(root), desc(X) = [L, R] + desc(R) + desc(L)
static bool number_nodes(const std::vector<GpuTopNode> &top, std::vector<float> &aabb, std::vector<int32_t> &links) {
    const size_t nt = top.size();
    std::vector<uint32_t> desc(nt, 0);
    for (size_t t = nt; t-- > 0;) {
        if (top[t].kind == rtb_emul::kTopHost) return false;
        if (top[t].kind == rtb_emul::kTopInternal) desc[t] = 2 + desc[(size_t) top[t].left] + desc[(size_t) top[t].right];
    }
    aabb.assign((size_t) (1 + desc[0]) * 6, 0.f);
    links.assign((size_t) (1 + desc[0]) * 5, 0);
    struct Item { uint32_t t, ref, base; };
    std::vector<Item> st{{0, 0, 1}};
    while (!st.empty()) {
        const Item it = st.back();
        st.pop_back();
        const GpuTopNode &g = top[it.t];
        std::memcpy(&aabb[(size_t) it.ref * 6], g.box, 24);
        int32_t *l = &links[(size_t) it.ref * 5];
        l[2] = g.split_dim;
        if (g.kind == rtb_emul::kTopLeaf) { l[0] = l[1] = -1; l[3] = (int32_t) g.first; l[4] = (int32_t) g.count; continue; }
        const uint32_t L = it.base, R = it.base + 1, r_base = it.base + 2, l_base = r_base + desc[(size_t) g.right];
        l[0] = (int32_t) L; l[1] = (int32_t) R; l[3] = -1; l[4] = 0;
        st.push_back({(uint32_t) g.left, L, l_base});
        st.push_back({(uint32_t) g.right, R, r_base});
    }
    return true;
}

static int check(const char *what, const std::vector<float> &pos, int depth_override, bool expect_handed_back = false) {
    const size_t n = pos.size() / 9;
    std::vector<float> tbox, ctr;
    boxes_of(pos, tbox, ctr);
    std::vector<GpuTopNode> top;
    std::vector<uint32_t> order;
    rtb_emul::GpuBuildStats gs;
    std::string err;
    if (!rtb::gpu_build_top_emulated(tbox.data(), ctr.data(), n, 5, 0, depth_override, top, order, gs, err)) {
        std::printf("FAILED %s: emulated builder: %s\n", what, err.c_str());
        return 1;
    }
    if (expect_handed_back) {   // a large node no cut improves goes back to the host uncut, in the order it arrived in
        bool ok = top.size() == 1 && top[0].kind == rtb_emul::kTopHost && top[0].first == 0 && top[0].count == n && gs.host_fallback_nodes == 1;
        for (size_t i = 0; ok && i < n; ++i) ok = order[i] == (uint32_t) i;
        std::printf(ok ? "ok  %-46s %8zu triangles: handed back to the host uncut, order restored\n" : "FAILED %s: expected the root handed back uncut (%zu)\n", what, n);
        return ok ? 0 : 1;
    }
    std::vector<float> aabb;
    std::vector<int32_t> links;
    if (!number_nodes(top, aabb, links)) { std::printf("FAILED %s: unexpected hand-over to the host\n", what); return 1; }
    if (depth_override >= 0) setenv("RTB_BUILD_SORT_DEPTH", std::to_string(depth_override).c_str(), 1);
    else unsetenv("RTB_BUILD_SORT_DEPTH");
    std::vector<uint32_t> h_order(n);
    std::vector<float> h_aabb((2 * n) * 6);
    std::vector<int32_t> h_links((2 * n) * 5);
    size_t h_nodes = 0;
    if (rtb_build_bvh(pos.data(), n, h_order.data(), h_aabb.data(), h_links.data(), 2 * n, &h_nodes, nullptr) != 0) {
        std::printf("FAILED %s: rtb_build_bvh: %s\n", what, rtb_last_error());
        return 1;
    }
    int bad = 0;
    if (h_nodes * 5 != links.size()) { std::printf("FAILED %s: %zu nodes, host %zu\n", what, links.size() / 5, h_nodes); return 1; }
    if (std::memcmp(order.data(), h_order.data(), n * 4) != 0) { std::printf("FAILED %s: triangle order differs\n", what); bad = 1; }
    if (std::memcmp(links.data(), h_links.data(), links.size() * 4) != 0) { std::printf("FAILED %s: links differ\n", what); bad = 1; }
    if (std::memcmp(aabb.data(), h_aabb.data(), aabb.size() * 4) != 0) { std::printf("FAILED %s: boxes differ (bits)\n", what); bad = 1; }
    if (!bad) std::printf("ok  %-46s %8zu triangles %8zu nodes  levels %2d rounds %4ld heaps %ld\n", what, n, h_nodes, gs.levels, gs.rounds, gs.heap_ranges);
    return bad;
}

int main() {
    setenv("RTB_BUILD_GPU", "0", 1);   // rtb_build_bvh: the host builder
    std::mt19937 g(2024);
    std::uniform_real_distribution<float> u(-1.f, 1.f);
    auto soup = [&](size_t n, float half, float q) {
        std::vector<float> p(n * 9);
        for (size_t i = 0; i < n; ++i) {
            float c[3] = {u(g) * half, u(g) * half, u(g) * half}, v[3][3];
            for (auto &w : v)
                for (int k = 0; k < 3; ++k) w[k] = c[k] + u(g) * 0.15f;
            if (q > 0.f)
                for (auto &w : v)
                    for (int k = 0; k < 3; ++k) w[k] = std::round(w[k] * q) / q;
            for (int k = 0; k < 3; ++k) { p[i * 9 + k] = v[0][k]; p[i * 9 + 3 + k] = v[1][k] - v[0][k]; p[i * 9 + 6 + k] = v[2][k] - v[0][k]; }
        }
        return p;
    };
    int bad = 0;
    for (size_t n : {5u, 6u, 9u, 17u, 33u, 100u, 1000u, 20000u}) bad |= check("random soup", soup(n, 3.f, 0.f), -1);
    bad |= check("random soup, 150 k", soup(150000, 5.3f, 0.f), -1);
    bad |= check("quantised soup (ties at every node)", soup(30000, 3.f, 4.f), -1);
    {
        std::vector<float> p = soup(8000, 2.f, 8.f), d = p;
        p.insert(p.end(), d.begin(), d.end());        // every triangle twice
        p.insert(p.end(), d.begin(), d.begin() + 9 * 3000);
        bad |= check("duplicated triangles", p, -1);
    }
    {
        // flat grid of axis-aligned right triangles in the plane z = 0, coordinates that include zeros of both signs
        std::vector<float> p;
        for (int y = -20; y < 20; ++y)
            for (int x = -20; x < 20; ++x)
                for (int h = 0; h < 2; ++h) {
                    const float x0 = (float) x * 0.25f, y0 = (float) y * 0.25f, z0 = (x + y) % 2 ? 0.f : -0.f;
                    const float a[9] = {x0, y0, z0, 0.25f, 0.f, 0.f, 0.f, 0.25f, 0.f}, b[9] = {x0 + 0.25f, y0 + 0.25f, z0, -0.25f, 0.f, 0.f, 0.f, -0.25f, 0.f};
                    p.insert(p.end(), h ? b : a, (h ? b : a) + 9);
                }
        bad |= check("flat axis-aligned grid, z = +-0", p, -1);
    }
    {
        const float t[9] = {0.5f, 0.25f, -1.f, 1.f, 0.f, 0.f, 0.f, 1.f, 0.5f};
        std::vector<float> p;
        for (int i = 0; i < 200; ++i) p.insert(p.end(), t, t + 9);
        bad |= check("200 coincident triangles (no cut improves)", p, -1);
        for (int i = 0; i < 800; ++i) p.insert(p.end(), t, t + 9);
        bad |= check("1000 coincident triangles (handed back)", p, -1, true);
        std::vector<float> q = soup(3000, 2.f, 0.f);
        for (int i = 0; i < 40; ++i) q.insert(q.begin() + 9 * 100, t, t + 9);
        for (int i = 0; i < 9; ++i) q.insert(q.end(), t, t + 9);
        bad |= check("coincident triangles inside a soup (big leaves)", q, -1);
    }
    {
        std::vector<float> p = soup(6000, 1.f, 0.f);
        for (size_t i = 0; i < p.size() / 9; ++i) p[i * 9] += (i % 2) ? 50.f : -50.f;     // two clusters far apart
        bad |= check("two distant clusters", p, -1);
        for (size_t i = 0; i < p.size() / 9; ++i) p[i * 9] = (float) i * 0.01f;            // centroids already sorted along x
        bad |= check("sorted along x", p, -1);
    }
    for (int d : {0, 2, 5}) bad |= check(d == 0 ? "depth limit 0 (everything heap-sorted)" : "depth limit cut (heap-sorted ranges)", soup(12000, 2.5f, d == 5 ? 6.f : 0.f), d);
    // ---- the hand-over between the two builders (host/bvh_builder.cpp: adopt_all) with the emulated builder in the GPU's place:
    // rtb_build_bvh under RTB_BUILD_GPU=1 and several hand-over sizes against the host build
    rtb::set_gpu_builder_for_tests(&rtb::gpu_build_top_emulated);
    unsetenv("RTB_BUILD_SORT_DEPTH");
    for (size_t n : {300u, 5000u, 60000u}) {
        const std::vector<float> pos = soup(n, 3.f, n == 5000u ? 4.f : 0.f);
        std::vector<uint32_t> o0(n), o1(n);
        std::vector<float> a0(2 * n * 6), a1(2 * n * 6);
        std::vector<int32_t> l0(2 * n * 5), l1(2 * n * 5);
        size_t n0 = 0, n1 = 0;
        setenv("RTB_BUILD_GPU", "0", 1);
        int rc = rtb_build_bvh(pos.data(), n, o0.data(), a0.data(), l0.data(), 2 * n, &n0, nullptr);
        for (const char *below : {"5", "64", "512"}) {
            setenv("RTB_BUILD_GPU", "1", 1);
            setenv("RTB_BUILD_GPU_BELOW", below, 1);
            double info[8] = {0};
            rc |= rtb_build_bvh(pos.data(), n, o1.data(), a1.data(), l1.data(), 2 * n, &n1, info);
            const bool used = info[0] == 1.0 || n < (size_t) std::atoi(below);
            const bool same = rc == 0 && n0 == n1 && std::memcmp(o0.data(), o1.data(), n * 4) == 0 && std::memcmp(l0.data(), l1.data(), n0 * 20) == 0 &&
                              std::memcmp(a0.data(), a1.data(), n0 * 24) == 0;
            std::printf("%s  hand-over below %-4s %38s %8zu triangles %8zu nodes\n", same && used ? "ok" : "FAILED", below, "", n, n1);
            if (!same || !used) bad = 1;
        }
    }
    rtb::set_gpu_builder_for_tests(nullptr);
    std::printf("%s\n", bad ? "FAILED" : "ok");
    return bad;
}
