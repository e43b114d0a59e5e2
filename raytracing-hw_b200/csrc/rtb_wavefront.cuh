// Wavefront (ray-queue) form of Scene::render (src/core/scene.cpp:17-65) -- per-pixel path state and the two
// per-slot steps (init, shade) that sit either side of the closest-hit queue (rtb_trace.cuh).
//
// Why (ncu, profiles/r01_k_render_lines.txt): the megakernel (rtb_integrator.cu: one thread runs a whole pixel)
// executes at 6.8 active threads per warp instruction -- traversal loops of different length, shading and the
// rejection loops all serialise inside a warp -- and is issue-bound.  Here a pixel is a SLOT of state in HBM; every
// iteration the slots' current rays go through the persistent closest-hit queue (lanes refill themselves, node and
// triangle phases are voted per warp), then one thread per slot shades its hit, draws the next direction from the
// slot's own sequential minstd_rand stream and either extends the path or folds it, accumulates and starts the
// pixel's next sample.  A slot never has more than one ray in flight, because sample s of a pixel cannot start
// before sample s-1 has consumed its part of the stream (src/core/scene.cpp:34-35).
//
// The per-slot arithmetic is the megakernel's (rtb_shade.cuh: same functions, same order), so both integrators give
// bit-identical sums.  Everything here is RTB_HD so that tests/devsim can drive the same step functions on the host.
#pragma once
#include "rtb_kernels.cuh"
#include "rtb_shade.cuh"

namespace rtb {

constexpr unsigned kSlotDone = 0x80000000u;   // high bit of an active-list entry: the slot has finished all samples
constexpr unsigned kSlotSuspended = 0x40000000u;   // the slot's ray was suspended by the queue kernel (rtb_trace.cuh: tail_steps): no hit yet, the shade step skips it
constexpr unsigned kSlotIndex = 0x3FFFFFFFu;
constexpr int kWfTileW = 8, kWfTileH = 4;     // a group of 32 consecutive slots is an 8x4 pixel tile

// Path state, SoA over slots (180 GB of HBM: ~0.3 kB per pixel is nothing; 1080p = 0.6 GB)
struct WfState {
    unsigned n_slots;        // tiles * 32
    int depth_cap;           // stack planes allocated (ray_depth - 1)
    float4 *ray_a;           // ox oy oz dx      (direction as handed to Ray::Ray, normalised again on load)
    float2 *ray_b;           // dy dz
    float4 *hit;             // t u v id(bits)   written by the queue, read by the shade step
    uint4 *ctl;              // x: minstd_rand state, y: normal_distribution::_M_saved (bits), z: sample index,
                             // w: bits 0-3 bounce k, 4-7 stack entries n, 8 _M_saved_available
    float4 *sum;             // per-pixel radiance sum (scene.cpp:42), w unused
    float4 *stk_e;           // [depth_cap][n_slots] emission.xyz, 1/pdf          (struct Bounce, rtb_shade.cuh)
    float4 *stk_m;           // [depth_cap][n_slots] material.xyz, cosine
    float *stk_a;            // [depth_cap][n_slots] material alpha
    int2 *susp;              // [n_slots][susp_stride] traversal state of a suspended ray (LaneTraversal::save3)
    unsigned susp_stride;    // kSuspHeader + BVH depth + 3 entries
};

RTB_HD bool wf_slot_pixel(const RenderLaunch &p, unsigned slot, int &px, int &lr, int &py) {
    const int tiles_x = (p.width + kWfTileW - 1) / kWfTileW;
    if (slot >= p.plane_slots) slot %= p.plane_slots;   // (stream planes beyond the first: same pixel layout)
    const unsigned tile = slot >> 5, lane = slot & 31u;
    const int tx = (int) (tile % (unsigned) tiles_x), ty = (int) (tile / (unsigned) tiles_x);
    px = tx * kWfTileW + (int) (lane & (kWfTileW - 1));
    lr = ty * kWfTileH + (int) (lane >> 3);
    py = p.row_begin + lr;
    if (p.il_count > 1) py = p.row_begin + ((lr / p.il_rows) * p.il_count + p.il_index) * p.il_rows + lr % p.il_rows;
    return px < p.width && lr < p.local_rows && py < p.row_end;
}

// sample range [s0, s1) of the stream a slot runs (rtb_render_params::substreams; one stream: [sample_begin, samples))
RTB_HD void wf_slot_samples(const RenderLaunch &p, unsigned slot, int &stream, int &s0, int &s1) {
    if (p.substreams <= 1) {
        stream = 0;
        s0 = p.sample_begin;
        s1 = p.samples;
        return;
    }
    stream = p.sub_begin + (int) (slot / p.plane_slots);
    s0 = (int) ((long long) stream * p.samples / p.substreams);
    s1 = (int) ((long long) (stream + 1) * p.samples / p.substreams);
}
// slot of pixel (px, local row lr) in stream plane `plane` of the launch
RTB_HD unsigned wf_pixel_slot(const RenderLaunch &p, int px, int lr, int plane) {
    const int tiles_x = (p.width + kWfTileW - 1) / kWfTileW;
    const unsigned tile = (unsigned) (lr / kWfTileH) * (unsigned) tiles_x + (unsigned) (px / kWfTileW);
    return (unsigned) plane * p.plane_slots + tile * 32u + (unsigned) ((lr % kWfTileH) * kWfTileW + px % kWfTileW);
}

RTB_HD void wf_store_ray(const WfState &w, unsigned slot, V3 o, V3 d) {
    w.ray_a[slot] = make_float4(o.x, o.y, o.z, d.x);
    w.ray_b[slot] = make_float2(d.y, d.z);
}
RTB_HD void wf_load_ray(const WfState &w, unsigned slot, Ray &r) {
    const float4 a = w.ray_a[slot];
    const float2 b = w.ray_b[slot];
    make_ray(r, a.x, a.y, a.z, a.w, b.x, b.y);
}

// first loop header of Scene::render for one pixel (scene.cpp:34-39): engine, sample 0's jitter and primary ray.
// Returns the active-list entry of the slot.
RTB_HD unsigned wf_init_slot(const DScene &sc, const RenderLaunch &p, const WfState &w, unsigned slot) {
    int px, lr, py;
    if (!wf_slot_pixel(p, slot, px, lr, py)) return slot | kSlotDone;
    Rng g;
    float4 sum0 = make_float4(0.f, 0.f, 0.f, 0.f);
    int stream, s0, s1;
    wf_slot_samples(p, slot, stream, s0, s1);
    if (s0 >= s1) {   // more streams than samples: this one has none (its sum is zero)
        w.sum[slot] = sum0;
        return slot | kSlotDone;
    }
    if (p.sample_begin > 0) {   // resume: the pixel's engine and sum as an earlier call left them
        const size_t o = ((size_t) lr * p.width + px) * 2;
        const uint4 a = p.d_pixel_state[o], b = p.d_pixel_state[o + 1];
        g.x = a.x;
        g.saved = u2f(a.y);
        g.has_saved = (a.z & 1u) != 0u;
        sum0 = make_float4(u2f(b.x), u2f(b.y), u2f(b.z), 0.f);
    } else {
        // stream k of a pixel is seeded like the pixel k canvases further down (stream 0: the reference's own seed)
        const uint64_t seed = (uint64_t) stream * (uint64_t) p.width * (uint64_t) p.height + (uint64_t) py * (uint64_t) p.width + (uint64_t) px;
        rng_seed(g, seed != 0 ? seed : (uint64_t) p.pixel0_seed);
    }
    const float ox = rng_pixel_offset(g);
    const float oy = rng_pixel_offset(g);
    // Camera::cast_in_pixel hands the un-normalised direction to Ray::Ray; store exactly that
    const V3 d = camera_dir(sc, px, py, p.width, p.height, ox, oy);
    wf_store_ray(w, slot, V3{sc.cam_pos[0], sc.cam_pos[1], sc.cam_pos[2]}, d);
    w.ctl[slot] = make_uint4(g.x, f2u(g.saved), (unsigned) s0, (g.has_saved ? 256u : 0u));
    w.sum[slot] = sum0;
    return slot;
}

// One vertex of Scene::intersect (scene.cpp:71-156) for the slot whose closest hit has just been written, plus --
// when the path ends here -- the fold of the recursion (scene.cpp:154), the accumulation (scene.cpp:42) and the
// next sample's primary ray (scene.cpp:36-39).  Returns the slot's next active-list entry.
template <bool COUNT>
RTB_HD unsigned wf_shade_slot(const DScene &sc, const RenderLaunch &p, const WfState &w, unsigned slot, PathCounters *pc) {
    const uint4 c = w.ctl[slot];
    Rng g;
    g.x = c.x;
    g.saved = u2f(c.y);
    g.has_saved = (c.w >> 8) & 1u;
    int sample = (int) c.z;
    int k = (int) (c.w & 15u), n = (int) ((c.w >> 4) & 15u);
    Ray r;
    wf_load_ray(w, slot, r);
    const float4 hv = w.hit[slot];
    Hit h;
    h.t = hv.x; h.u = hv.y; h.v = hv.z; h.id = (int) f2u(hv.w);
    const int depth = sc.ray_depth;

    bool ended = true;
    V3 last = V3{0.f, 0.f, 0.f};      // colour of the deepest successful vertex when the path ends at a hit
    bool last_is_hit = false;
    if (h.id >= 0 && h.t < sc.max_distance) {
        const float4 q2 = ldg(sc.bvh.tris + (size_t) h.id * kTriStride + 2);
        V3 Ng = V3{q2.y, q2.z, q2.w};
        const V3 rd = V3{r.d[0], r.d[1], r.d[2]};
        const bool inside = dot(rd, Ng) > 0.f;   // primitive.cpp:51-54
        if (inside) Ng = -Ng;
        const Shade s = load_shade(sc, h.id);
        const Material m = load_material(sc, s.mesh);
        const V3 emission = textured_factor(sc, s, m.emission_i, -1, m.emission, h.u, h.v);
        last = emission;
        last_is_hit = true;
        const V3 pos = V3{r.o[0], r.o[1], r.o[2]} + rd * h.t;
        V3 Ns = shading_normal(sc, h.id, s, m, h.u, h.v);
        if (inside) Ns = -Ns;
        float rough2, metallic;
        metallic_roughness(sc, s, m, h.u, h.v, rough2, metallic);
        rough2 = std_max(kRoughness2Limit, rough2);
        const V3 eye = -rd;
        const V3 dir = sample_direction(sc, pos, Ns, eye, rough2, g);   // consumes the stream even at the last vertex
        bool go = k + 1 < depth;
        if (go && dot(dir, Ns) <= 0.f) {
            if (dot(dir, Ng) <= 0.f) go = false;
            else Ns = Ng;
        }
        if (go) {
            const float pdf = pdf_direction<COUNT>(sc, pos, Ns, eye, rough2, dir, COUNT ? &pc->light : nullptr, &pc->light_queries);
            if (!(pdf <= 0.f || pdf != pdf)) {
                const V3 half = normal(dir - rd);
                const float vis = smith_g(rough2, Ns, eye, dir) * (1.f / (4.f * fabsf(dot(Ns, rd)) * fabsf(dot(Ns, dir))));
                const float spec = ggx_d(rough2, Ns, half) * vis;
                const float VdotH = fabsf(dot(eye, half));
                const V3 base = textured_factor(sc, s, m.base_i, m.fused, m.base, h.u, h.v);
                const float p5 = pow5_via_double(1.f - VdotH);
                const V3 metal = spec * V3{fresnel(base.x, p5), fresnel(base.y, p5), fresnel(base.z, p5)};
                const V3 diffuse = base * kInvPi;
                const float dsc = fresnel(0.04f, p5);
                const V3 dielectric = diffuse * (1.f - dsc) + (V3{1.f, 1.f, 1.f} * spec) * dsc;
                const V3 material = dielectric * (1.f - metallic) + metal * metallic;
                const size_t e = (size_t) n * w.n_slots + slot;
                w.stk_e[e] = make_float4(emission.x, emission.y, emission.z, 1.f / pdf);
                w.stk_m[e] = make_float4(material.x, material.y, material.z, dot(dir, Ns));
                w.stk_a[e] = m.alpha;
                ++n;
                ++k;
                wf_store_ray(w, slot, pos + dir * kStep, dir);   // Ray reflect_ray(pos + dir * step, dir)
                ended = false;
            }
        }
    }

    unsigned entry = slot;
    if (ended) {
        // fold the recursion back to front: L_j = E_j + ((((L_{j+1} * coeff) * material) * cosine) * alpha)
        V3 L;
        int j;
        if (last_is_hit) {
            L = last;
            j = n - 1;
        } else if (n == 0) {
            L = V3{sc.bg_color[0], sc.bg_color[1], sc.bg_color[2]};
            j = -1;
        } else {   // the deepest stored vertex had no successful child: it contributes its emission only
            const float4 e = w.stk_e[(size_t) (n - 1) * w.n_slots + slot];
            L = V3{e.x, e.y, e.z};
            j = n - 2;
        }
        for (; j >= 0; --j) {
            const size_t o = (size_t) j * w.n_slots + slot;
            const float4 e = w.stk_e[o], mt = w.stk_m[o];
            const float a = w.stk_a[o];
            L = V3{e.x, e.y, e.z} + (((L * e.w) * V3{mt.x, mt.y, mt.z}) * mt.w) * a;
        }
        const float4 s4 = w.sum[slot];
        const V3 sum = V3{s4.x, s4.y, s4.z} + L;
        ++sample;
        k = 0;
        n = 0;
        int px, lr, py;
        wf_slot_pixel(p, slot, px, lr, py);
        int stream, s0, s1;
        wf_slot_samples(p, slot, stream, s0, s1);
        if (sample >= s1 && p.substreams > 1) {
            // one of several streams of the pixel: its sum waits in the slot for wf_fold_pixel
            w.sum[slot] = make_float4(sum.x, sum.y, sum.z, 0.f);
            entry = slot | kSlotDone;
        } else if (sample >= s1) {
            const size_t o = ((size_t) lr * p.width + px) * 3;
            if (p.d_accum) {
                p.d_accum[o] = sum.x;
                p.d_accum[o + 1] = sum.y;
                p.d_accum[o + 2] = sum.z;
            }
            if (p.d_rgb8) {
                const float normalizer = 1.f / (float) p.samples;
                p.d_rgb8[o] = resolve_channel(sum.x, normalizer);
                p.d_rgb8[o + 1] = resolve_channel(sum.y, normalizer);
                p.d_rgb8[o + 2] = resolve_channel(sum.z, normalizer);
            }
            if (p.d_pixel_state) {
                const size_t so = ((size_t) lr * p.width + px) * 2;
                p.d_pixel_state[so] = make_uint4(g.x, f2u(g.saved), g.has_saved ? 1u : 0u, (unsigned) sample);
                p.d_pixel_state[so + 1] = make_uint4(f2u(sum.x), f2u(sum.y), f2u(sum.z), 0u);
            }
            entry = slot | kSlotDone;
        } else {
            w.sum[slot] = make_float4(sum.x, sum.y, sum.z, 0.f);
            const float ox = rng_pixel_offset(g);
            const float oy = rng_pixel_offset(g);
            const V3 d = camera_dir(sc, px, py, p.width, p.height, ox, oy);
            wf_store_ray(w, slot, V3{sc.cam_pos[0], sc.cam_pos[1], sc.cam_pos[2]}, d);
        }
    }
    w.ctl[slot] = make_uint4(g.x, f2u(g.saved), (unsigned) sample, (unsigned) k | ((unsigned) n << 4) | (g.has_saved ? 256u : 0u));
    return entry;
}

// Sample-range split: the sums of the launch's streams of pixel (px, local row lr), folded in stream order
// (src/core/scene.cpp:42 accumulates a pixel's samples in order; here each stream's samples are in order and the streams
// follow one another), written to d_accum and -- when the launch holds every stream of the pixel -- resolved to d_rgb8.
RTB_HD void wf_fold_pixel(const RenderLaunch &p, const WfState &w, int px, int lr) {
    const int planes = p.sub_end - p.sub_begin;
    const float4 s = w.sum[wf_pixel_slot(p, px, lr, 0)];
    V3 sum = V3{s.x, s.y, s.z};
    for (int k = 1; k < planes; ++k) {
        const float4 t = w.sum[wf_pixel_slot(p, px, lr, k)];
        sum = sum + V3{t.x, t.y, t.z};
    }
    const size_t o = ((size_t) lr * p.width + px) * 3;
    if (p.d_accum) {
        p.d_accum[o] = sum.x;
        p.d_accum[o + 1] = sum.y;
        p.d_accum[o + 2] = sum.z;
    }
    if (p.d_rgb8) {
        const float normalizer = 1.f / (float) p.samples;
        p.d_rgb8[o] = resolve_channel(sum.x, normalizer);
        p.d_rgb8[o + 1] = resolve_channel(sum.y, normalizer);
        p.d_rgb8[o + 2] = resolve_channel(sum.z, normalizer);
    }
}

// ---- launchers (rtb_wavefront.cu)
size_t wf_state_bytes(unsigned n_slots, int depth_cap, int bvh_depth);

}  // namespace rtb
