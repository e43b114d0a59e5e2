// CPU-side debugging harness: the device functions of raytracing-hw_b200/csrc/*.cuh compiled for the HOST by nvcc,
// driven over host arrays.  Lets the decision-exact restatement be checked against the reference in a container
// without a GPU (tests/test_devsim.py).  TEST TOOLING ONLY: librtb.so neither links nor calls this, and nothing
// here is a fallback for the C ABI (which fails with RTB_ERR_NO_DEVICE when no GPU is usable).
#include "../../raytracing-hw_b200/csrc/rtb_layout.hpp"
#include "../../raytracing-hw_b200/csrc/rtb_wavefront.cuh"
#include "../../raytracing-hw_b200/csrc/rtb_trace.cuh"
#include "../../raytracing-hw_b200/csrc/rtb_assemble.cuh"

#include <omp.h>

using namespace rtb;

namespace {
struct SimScene {
    HostLayout L;
    std::vector<DTexture> tex;
    std::vector<std::vector<uint4>> quads, fused_data;
    std::vector<DFused> fused;
    DScene d;
};

bool make_scene(const rtb_scene_desc *desc, SimScene &s, std::string &err) {
    std::memset(&s.d, 0, sizeof(s.d));
    if (!build_layout(desc, s.L, err)) return false;
    auto bind = [](HostBvhOut &h, DBvh &d, uint32_t n) {
        d.nodes = h.nodes.data();
        d.tris = h.tris.data();
        d.big_leaves = h.big.data();
        fill_bvh_scalars(h, d, n);
    };
    bind(s.L.bvh, s.d.bvh, desc->n_tris);
    if (desc->n_lights) {
        bind(s.L.light_bvh, s.d.light_bvh, desc->n_lights);
        s.d.light_inv_area = s.L.light_bvh.inv_area.data();
    }
    s.d.tri_shade = s.L.shade.data();
    s.d.tri_tan = s.L.tan.empty() ? nullptr : s.L.tan.data();
    s.d.mesh_material = s.L.material.data();
    s.d.mesh_normal_transform = s.L.normal_transform.data();
    s.d.n_meshes = s.L.n_meshes;
    s.tex.resize(desc->n_textures);
    s.quads.resize(desc->n_textures);
    for (uint32_t i = 0; i < desc->n_textures; ++i) {
        const int w = desc->textures[i].width, h = desc->textures[i].height;
        s.quads[i].resize((size_t) w * h);
        for (int y = 0; y < h; ++y)
            for (int x = 0; x < w; ++x) s.quads[i][(size_t) y * w + x] = texture_quad((const uchar4 *) desc->textures[i].rgba, w, h, x, y);
        s.tex[i] = {(const uchar4 *) desc->textures[i].rgba, s.quads[i].data(), w, h};
    }
    s.d.textures = s.tex.data();
    s.fused.resize(s.L.fused.size());
    s.fused_data.resize(s.L.fused.size());
    for (size_t g = 0; g < s.L.fused.size(); ++g) {
        const auto &G = s.L.fused[g];
        auto px = [&](int ti) { return ti >= 0 ? (const uchar4 *) desc->textures[ti].rgba : nullptr; };
        s.fused_data[g].resize((size_t) G.width * G.height * 4);
        for (int y = 0; y < G.height; ++y)
            for (int x = 0; x < G.width; ++x) fused_texel(px(G.tex[0]), px(G.tex[1]), px(G.tex[2]), G.width, G.height, x, y, &s.fused_data[g][((size_t) y * G.width + x) * 4]);
        s.fused[g] = {s.fused_data[g].data(), G.width, G.height};
    }
    s.d.fused = s.fused.data();
    fill_scalars(desc, s.d);
    fill_lut(s.L.lut);
    s.d.lut = s.L.lut.data();
    return true;
}
std::string g_err;
}  // namespace

// Form 3 of the per-lane state machine with a RING-entry ring on top of the stack; every > 0: the state goes through
// save3 / load3 into a FRESH lane (new stack, new ring, ray re-made from the ray record) every `every` steps, as the
// queue kernel does across launches.
template <int RING>
static Hit lane_form3(const DBvh &bvh, const float *ray6, int every, Counters &c) {
    using Lane = LaneTraversal<true, RING>;
    int2 stack2[Lane::kStack3], stack3[Lane::kStack3];
    int2 rec[Lane::kSuspHeader + kMaxStack + 3];
    Lane L;
    L.reset3(stack2);
    make_ray(L.r, ray6[0], ray6[1], ray6[2], ray6[3], ray6[4], ray6[5]);
    L.begin(bvh, &c);
    int left = every;
    while (!L.finished()) {
        L.step3(bvh, &c);
        if (every > 0 && --left <= 0 && !L.finished()) {
            L.save3(rec);
            Lane M;
            M.reset3(L.s2 == stack2 ? stack3 : stack2);
            make_ray(M.r, ray6[0], ray6[1], ray6[2], ray6[3], ray6[4], ray6[5]);
            M.load3(rec);
            L = M;
            left = every;
        }
    }
    L.finish3();
    return L.best;
}

extern "C" {

const char *devsim_last_error() { return g_err.c_str(); }

int devsim_closest_hit(const rtb_scene_desc *desc, const float *rays6, size_t n, int cull, int32_t *id, float *t, float *u, float *v,
                       uint8_t *inside, unsigned long long *counters2) {
    SimScene s;
    if (!make_scene(desc, s, g_err)) return 1;
    unsigned long long box = 0, tri = 0;
#pragma omp parallel for schedule(dynamic, 1024) reduction(+ : box, tri)
    for (size_t i = 0; i < n; ++i) {
        Ray r;
        make_ray(r, rays6[i * 6], rays6[i * 6 + 1], rays6[i * 6 + 2], rays6[i * 6 + 3], rays6[i * 6 + 4], rays6[i * 6 + 5]);
        Hit h;
        Counters c = {0, 0};
        const bool ok = cull == 0 ? closest_hit<0, true>(s.d.bvh, r, h, &c) : closest_hit<1, true>(s.d.bvh, r, h, &c);
        box += c.box;
        tri += c.tri;
        id[i] = ok ? h.id : -1;
        if (t) t[i] = ok ? h.t : 0.f;
        if (u) u[i] = ok ? h.u : 0.f;
        if (v) v[i] = ok ? h.v : 0.f;
        if (inside) {
            uint8_t in = 0;
            if (ok) {
                const float4 q2 = s.d.bvh.tris[(size_t) h.id * kTriStride + 2];
                in = dot(V3{r.d[0], r.d[1], r.d[2]}, V3{q2.y, q2.z, q2.w}) > 0.f ? 1 : 0;
            }
            inside[i] = in;
        }
    }
    if (counters2) { counters2[0] = box; counters2[1] = tri; }
    return 0;
}

// The queue kernels' per-lane state machine (rtb_trace.cuh: LaneTraversal, bounded branch-free pops) on the host.
// schedule_seed selects the form: 0..2 the first form; 3 form 3 (one pop site per step); 4 / 5 form 3 with a 2- / 8-entry
// ring on top of the stack; 6 form 3 with the top entry in registers; 100 + n form 3 suspended every n steps; 10000 + n
// the same with the 2-entry ring; 20000 + n the same with the top entry in registers.
int devsim_closest_hit_lane(const rtb_scene_desc *desc, const float *rays6, size_t n, unsigned schedule_seed, int32_t *id, float *t,
                            float *u, float *v, unsigned long long *counters2) {
    SimScene s;
    if (!make_scene(desc, s, g_err)) return 1;
    unsigned long long box = 0, tri = 0;
#pragma omp parallel for schedule(dynamic, 1024) reduction(+ : box, tri)
    for (size_t i = 0; i < n; ++i) {
        Counters c = {0, 0};
        Hit best;
        const float *ray6 = rays6 + i * 6;
        if (schedule_seed >= 20000u) best = lane_form3<-1>(s.d.bvh, ray6, (int) (schedule_seed - 20000u), c);
        else if (schedule_seed >= 10000u) best = lane_form3<2>(s.d.bvh, ray6, (int) (schedule_seed - 10000u), c);
        else if (schedule_seed >= 100u) best = lane_form3<0>(s.d.bvh, ray6, (int) (schedule_seed - 100u), c);
        else if (schedule_seed == 6u) best = lane_form3<-1>(s.d.bvh, ray6, 0, c);
        else if (schedule_seed == 5u) best = lane_form3<8>(s.d.bvh, ray6, 0, c);
        else if (schedule_seed == 4u) best = lane_form3<2>(s.d.bvh, ray6, 0, c);
        else if (schedule_seed == 3u) best = lane_form3<0>(s.d.bvh, ray6, 0, c);
        else {
            int stack_link[kMaxStack + 1];
            float stack_m[kMaxStack + 1];
            LaneTraversal<true> L;
            L.reset(stack_link, stack_m);
            make_ray(L.r, ray6[0], ray6[1], ray6[2], ray6[3], ray6[4], ray6[5]);
            L.begin(s.d.bvh, &c);
            while (!L.finished()) L.step(s.d.bvh, &c);
            best = L.best;
        }
        box += c.box;
        tri += c.tri;
        const bool ok = best.id >= 0;
        id[i] = ok ? best.id : -1;
        if (t) t[i] = ok ? best.t : 0.f;
        if (u) u[i] = ok ? best.u : 0.f;
        if (v) v[i] = ok ? best.v : 0.f;
    }
    if (counters2) { counters2[0] = box; counters2[1] = tri; }
    return 0;
}

int devsim_light_pdf(const rtb_scene_desc *desc, const float *rays6, size_t n, float *pdf) {
    SimScene s;
    if (!make_scene(desc, s, g_err)) return 1;
    for (size_t i = 0; i < n; ++i) {
        const float *rp = rays6 + i * 6;
        pdf[i] = s.d.n_lights ? pdf_lights<false>(s.d, V3{rp[0], rp[1], rp[2]}, V3{rp[3], rp[4], rp[5]}, nullptr) : 0.f;
    }
    return 0;
}

int devsim_render(const rtb_scene_desc *desc, const rtb_render_params *p, float *accum, uint8_t *rgb8, unsigned long long *counters4) {
    SimScene s;
    if (!make_scene(desc, s, g_err)) return 1;
    const int cull = (p->flags & RTB_RENDER_REFERENCE_CULL) ? 0 : 1;
    unsigned long long c0 = 0, c1 = 0, c2 = 0, c3 = 0;
    const int rows = p->row_end - p->row_begin;
#pragma omp parallel for schedule(dynamic, 1) reduction(+ : c0, c1, c2, c3)
    for (int lr = 0; lr < rows; ++lr) {
        const int py = p->row_begin + lr;
        for (int px = 0; px < p->width; ++px) {
            PathCounters pc = {0, 0, {0, 0}, {0, 0}};
            const V3 sum = cull == 0 ? render_pixel<0, true>(s.d, px, py, p->width, p->height, p->samples, p->pixel0_seed, &pc)
                                     : render_pixel<1, true>(s.d, px, py, p->width, p->height, p->samples, p->pixel0_seed, &pc);
            c0 += pc.rays; c1 += pc.light_queries; c2 += pc.tests.box + pc.light.box; c3 += pc.tests.tri + pc.light.tri;
            const size_t o = ((size_t) lr * p->width + px) * 3;
            if (accum) { accum[o] = sum.x; accum[o + 1] = sum.y; accum[o + 2] = sum.z; }
            if (rgb8) {
                const float nz = 1.f / (float) p->samples;
                rgb8[o] = resolve_channel(sum.x, nz);
                rgb8[o + 1] = resolve_channel(sum.y, nz);
                rgb8[o + 2] = resolve_channel(sum.z, nz);
            }
        }
    }
    if (counters4) { counters4[0] = c0; counters4[1] = c1; counters4[2] = c2; counters4[3] = c3; }
    return 0;
}

// The wavefront integrator's per-slot steps (rtb_wavefront.cuh: wf_init_slot / wf_shade_slot -- the same functions the
// k_wf_* kernels call) driven on host arrays, with the queue replaced by closest_hit_global per active slot.
int devsim_render_wavefront_resume(const rtb_scene_desc *desc, const rtb_render_params *p, void *pixel_state, float *accum, uint8_t *rgb8,
                                   unsigned long long *counters4);
int devsim_render_wavefront(const rtb_scene_desc *desc, const rtb_render_params *p, float *accum, uint8_t *rgb8, unsigned long long *counters4) {
    return devsim_render_wavefront_resume(desc, p, nullptr, accum, rgb8, counters4);
}
// samples [p->sample_begin, p->samples) with the per-pixel state of rtb_render_resume (32 bytes per local pixel; in when
// sample_begin > 0, out when not null)
int devsim_render_wavefront_resume(const rtb_scene_desc *desc, const rtb_render_params *p, void *pixel_state, float *accum, uint8_t *rgb8,
                                   unsigned long long *counters4) {
    SimScene s;
    if (!make_scene(desc, s, g_err)) return 1;
    RenderLaunch L;
    std::memset(&L, 0, sizeof(L));
    L.width = p->width; L.height = p->height; L.samples = p->samples;
    L.row_begin = p->row_begin; L.row_end = p->row_end;
    L.il_count = p->interleave_count; L.il_index = p->interleave_index; L.il_rows = p->interleave_rows > 0 ? p->interleave_rows : 4;
    {   // rows this call renders (same rule as rtb_render_local_rows)
        const int rows = p->row_end - p->row_begin;
        L.local_rows = rows;
        if (p->interleave_count > 1) {
            const int R = L.il_rows, blocks = (rows + R - 1) / R;
            L.local_rows = 0;
            for (int b = p->interleave_index; b < blocks; b += p->interleave_count) L.local_rows += std::min(R, rows - b * R);
        }
    }
    L.pixel0_seed = p->pixel0_seed;
    L.d_accum = accum; L.d_rgb8 = rgb8;
    L.sample_begin = p->sample_begin;
    L.d_pixel_state = (uint4 *) pixel_state;
    if (p->sample_begin > 0 && !pixel_state) { g_err = "sample_begin > 0 without a pixel state"; return 3; }
    // sample-range split (rtb_render_params::substreams), as rtb_render_resume_device sets it up
    L.substreams = std::max(1, p->substreams);
    L.sub_begin = 0; L.sub_end = L.substreams;
    if (L.substreams > 1 && (p->substream_begin != 0 || p->substream_end != 0)) { L.sub_begin = p->substream_begin; L.sub_end = p->substream_end; }
    if (L.sub_begin < 0 || L.sub_begin >= L.sub_end || L.sub_end > L.substreams) { g_err = "bad substream range"; return 3; }
    L.plane_slots = (unsigned) (((L.width + kWfTileW - 1) / kWfTileW) * ((L.local_rows + kWfTileH - 1) / kWfTileH) * 32);
    const unsigned n_slots = L.plane_slots * (unsigned) (L.sub_end - L.sub_begin);
    const int cap = std::max(1, s.d.ray_depth - 1);
    std::vector<float4> ray_a(n_slots), hit(n_slots), sum(n_slots), stk_e((size_t) n_slots * cap), stk_m((size_t) n_slots * cap);
    std::vector<float2> ray_b(n_slots);
    std::vector<uint4> ctl(n_slots);
    std::vector<float> stk_a((size_t) n_slots * cap);
    WfState w;
    w.n_slots = n_slots; w.depth_cap = cap;
    w.ray_a = ray_a.data(); w.ray_b = ray_b.data(); w.hit = hit.data(); w.ctl = ctl.data(); w.sum = sum.data();
    w.stk_e = stk_e.data(); w.stk_m = stk_m.data(); w.stk_a = stk_a.data();
    std::vector<unsigned> list(n_slots);
    for (unsigned i = 0; i < n_slots; ++i) list[i] = wf_init_slot(s.d, L, w, i);
    unsigned long long c0 = 0, c1 = 0, c2 = 0, c3 = 0;
    const long long chain = L.substreams > 1 ? ((long long) p->samples + L.substreams - 1) / L.substreams : (long long) (p->samples - p->sample_begin);
    const long long iterations = chain * s.d.ray_depth;
    for (long long it = 0; it < iterations; ++it) {
#pragma omp parallel for schedule(dynamic, 64) reduction(+ : c0, c1, c2, c3)
        for (long long i = 0; i < (long long) n_slots; ++i) {
            const unsigned e = list[(size_t) i];
            if (e & kSlotDone) continue;
            Ray r;
            wf_load_ray(w, e, r);
            Hit h;
            Counters c = {0, 0};
            closest_hit_global<true>(s.d.bvh, r, h, &c);
            w.hit[e] = make_float4(h.t, h.u, h.v, u2f((uint32_t) h.id));
            PathCounters pc = {0, 0, {0, 0}, {0, 0}};
            list[(size_t) i] = wf_shade_slot<true>(s.d, L, w, e, &pc);
            c0 += 1; c1 += pc.light_queries; c2 += c.box + pc.light.box; c3 += c.tri + pc.light.tri;
        }
    }
    for (unsigned i = 0; i < n_slots; ++i) {
        int px, lr, py;
        if (wf_slot_pixel(L, i, px, lr, py) && !(list[i] & kSlotDone)) { g_err = "a slot did not finish"; return 2; }
    }
    if (L.substreams > 1)
        for (int lr = 0; lr < L.local_rows; ++lr)
            for (int px = 0; px < L.width; ++px) wf_fold_pixel(L, w, px, lr);
    if (counters4) { counters4[0] = c0; counters4[1] = c1; counters4[2] = c2; counters4[3] = c3; }
    return 0;
}

// one path for an explicit ray and engine seed (mirror of oracle/ref_harness.cpp: ref_trace_path)
int devsim_trace_path(const rtb_scene_desc *desc, const float *o3, const float *d3, unsigned engine_seed, float *rgb3, unsigned *engine_after) {
    SimScene s;
    if (!make_scene(desc, s, g_err)) return 1;
    Rng g;
    rng_seed(g, engine_seed);
    Ray r;
    make_ray(r, o3[0], o3[1], o3[2], d3[0], d3[1], d3[2]);
    PathCounters pc = {0, 0, {0, 0}, {0, 0}};
    const V3 c = trace_path<0, false>(s.d, r, g, &pc);
    rgb3[0] = c.x; rgb3[1] = c.y; rgb3[2] = c.z;
    *engine_after = g.x;
    return 0;
}

// ---- scene assembly.  devsim_layout: the buffers rtb_scene_create uploads (rtb_layout.hpp, from the host-assembled arrays);
// devsim_assemble: the buffers rtb_scene_create_raw's kernels write, produced by running rtb_assemble.cuh's per-triangle /
// per-node functions on the host over the caller's raw views.  Buffers are caller-allocated: tris n x 16 floats, shade
// n x 16 floats, tan n x 12 floats (or null), nodes (n_nodes - 1) / 2 x 64 bytes, big `big_cap` int2; scalars =
// {root_link, n_internal, max_depth, n_big, has_tan}.
int devsim_layout(const rtb_scene_desc *desc, float *tris, float *shade, float *tan, void *nodes, int *big, int big_cap, int *scalars) {
    HostLayout L;
    if (!build_layout(desc, L, g_err)) return 1;
    std::memcpy(tris, L.bvh.tris.data(), L.bvh.tris.size() * sizeof(float4));
    std::memcpy(shade, L.shade.data(), L.shade.size() * sizeof(float4));
    if (tan && !L.tan.empty()) std::memcpy(tan, L.tan.data(), L.tan.size() * sizeof(float4));
    std::memcpy(nodes, L.bvh.nodes.data(), L.bvh.nodes.size() * sizeof(DNode));
    if ((int) L.bvh.big.size() > big_cap) { g_err = "big-leaf table larger than the caller's buffer"; return 1; }
    std::memcpy(big, L.bvh.big.data(), L.bvh.big.size() * sizeof(int2));
    scalars[0] = L.bvh.root_link; scalars[1] = (int) L.bvh.nodes.size(); scalars[2] = L.bvh.max_depth; scalars[3] = (int) L.bvh.big.size();
    scalars[4] = L.tan.empty() ? 0 : 1;
    return 0;
}

int devsim_assemble(const rtb_scene_desc *desc, const rtb_raw_geometry *raw, float *tris, float *shade, float *tan, void *nodes, int *big, int big_cap,
                    int *scalars) {
    HostLayout L;
    if (!build_layout_common(desc, L, g_err)) return 1;
    std::vector<DRawMesh> mesh(raw->n_meshes);
    std::vector<uint32_t> first(raw->n_meshes + 1, 0);
    for (uint32_t m = 0; m < raw->n_meshes; ++m) {
        const rtb_raw_mesh &r = raw->meshes[m];
        DRawMesh &d = mesh[m];
        std::memset(&d, 0, sizeof(d));
        first[m + 1] = first[m] + r.n_tris;
        if (r.n_tris == 0) continue;
        d.idx = (const uint8_t *) r.indices; d.pos = (const uint8_t *) r.position; d.nrm = (const uint8_t *) r.normal;
        d.tan = (const uint8_t *) r.tangent; d.uv = (const uint8_t *) r.texcoord;
        d.first_tri = first[m]; d.n_tris = r.n_tris; d.n_vertices = r.n_vertices;
        d.index_bytes = r.index_bytes; d.idx_stride = r.index_stride;
        d.pos_stride = r.position_stride; d.nrm_stride = r.normal_stride; d.tan_stride = r.tangent_stride; d.uv_stride = r.texcoord_stride;
        d.normal_mode = r.normal_mode; d.flags = r.flags; d.mesh_id = r.mesh_id;
        std::memcpy(d.transform, r.transform, sizeof(d.transform));
        std::memcpy(d.normal_transform, &L.normal_transform[(size_t) r.mesh_id * 16], sizeof(d.normal_transform));
    }
    int error = kAsmOk;
    const bool want_tan = L.any_normal_map && tan;
    for (uint32_t i = 0; i < desc->n_tris; ++i) {
        const uint32_t src = raw->tri_source[i];
        if (src >= first[raw->n_meshes]) { error = kAsmSource; break; }
        const int m = mesh_of_source(first.data(), (int) raw->n_meshes, src);
        const int e = assemble_triangle_records(mesh[(size_t) m], src - first[(size_t) m], (float4 *) tris + (size_t) i * kTriStride, (float4 *) shade + (size_t) i * 4,
                                                want_tan ? (float4 *) tan + (size_t) i * 3 : nullptr);
        if (e != kAsmOk && error == kAsmOk) error = e;
    }
    LeafLinker leaf;
    big_leaf_slots(desc->n_nodes, desc->node_links, desc->n_tris, L.bvh.big, leaf.big_node, leaf.big_slot);
    if ((int) L.bvh.big.size() > big_cap) { g_err = "big-leaf table larger than the caller's buffer"; return 1; }
    std::memcpy(big, L.bvh.big.data(), L.bvh.big.size() * sizeof(int2));
    const int32_t *root = desc->node_links;
    const bool root_leaf = root[4] > 0;
    const uint32_t n_internal = root_leaf ? 0u : (desc->n_nodes - 1u) / 2u;
    int levels = 0;
    if (!root_leaf) {
        const BigLeafIndex bigx{leaf.big_node.data(), leaf.big_slot.data(), (int) leaf.big_node.size()};
        std::vector<NodeItem> cur{NodeItem{0u, 0u, 1u, desc->n_nodes}}, next;
        while (!cur.empty() && error == kAsmOk) {
            ++levels;
            next.clear();
            for (const NodeItem &it : cur) {
                NodeItem kids[2];
                int nk = 0;
                const int e = convert_node(it, desc->n_nodes, n_internal, desc->n_tris, desc->node_aabb, desc->node_links, bigx, (DNode *) nodes, kids, &nk);
                if (e != kAsmOk) { error = e; break; }
                for (int k = 0; k < nk; ++k) next.push_back(kids[k]);
                if (next.size() > n_internal) { error = kAsmNodeOrder; break; }
            }
            cur.swap(next);
        }
    }
    if (error != kAsmOk) { g_err = assemble_error_text(error); return 2; }
    scalars[0] = root_leaf ? leaf.link(0, root[3], root[4]) : 0; scalars[1] = (int) n_internal; scalars[2] = levels; scalars[3] = (int) L.bvh.big.size();
    scalars[4] = want_tan ? 1 : 0;
    return 0;
}

}  // extern "C"
