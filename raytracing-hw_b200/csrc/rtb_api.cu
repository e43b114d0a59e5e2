// Device half of the C ABI (include/rtb.h): scene upload (reference-order nodes -> 64-byte two-child nodes),
// batched closest hit, light pdf, render.  No CPU fallback: every entry point needs a CUDA device.
#include "../../include/rtb.h"
#include "rtb_kernels.cuh"
#include "rtb_block_cache.h"
#include "rtb_layout.hpp"
#include "rtb_assemble.cuh"

#include <algorithm>
#include <cmath>
#include <cstring>
#include <string>
#include <mutex>
#include <thread>
#include <vector>

namespace rtb {
extern thread_local std::string g_last_error;   // host_api.cpp
void set_error(const std::string &e);
}  // namespace rtb

#define RTB_CUDA(call)                                                                                              \
    do {                                                                                                            \
        cudaError_t e_ = (call);                                                                                    \
        if (e_ != cudaSuccess) {                                                                                    \
            rtb::set_error(std::string(#call) + ": " + cudaGetErrorString(e_));                                     \
            return (e_ == cudaErrorNoDevice || e_ == cudaErrorInsufficientDriver) ? RTB_ERR_NO_DEVICE : RTB_ERR_CUDA; \
        }                                                                                                           \
    } while (0)

constexpr unsigned kQueueHeads = 64;

namespace rtb {
namespace {
BlockCache g_blocks;
}  // namespace
}  // namespace rtb

struct rtb_scene {
    int device = 0;
    rtb::DScene d;
    std::vector<std::pair<void *, size_t>> allocs;   // blocks from rtb::g_blocks
    size_t traversal_bytes = 0;
    size_t upload_bytes = 0;   // host->device bytes copied by rtb_scene_create / rtb_scene_create_raw
    size_t n_big = 0;          // entries of d.bvh.big_leaves
    unsigned *queue_heads = nullptr;   // kQueueHeads ray-queue cursors, rotated per launch (stream-ordered memset)
    unsigned next_head = 0;
    cudaStream_t streams[3] = {nullptr, nullptr, nullptr};
    cudaEvent_t ev[2] = {nullptr, nullptr};
    // staging buffers reused across calls
    void *stage = nullptr;
    size_t stage_bytes = 0;
    // path state of the wavefront integrator (rtb_wavefront.cuh), grown on demand
    void *wf_state = nullptr;
    size_t wf_bytes = 0;
    // buffers of the multi-GPU combine (rtb_dist.cu: local partial output, root staging), grown on demand
    void *scratch[3] = {nullptr, nullptr, nullptr};
    size_t scratch_bytes[3] = {0, 0, 0};
    unsigned long long launches = 0;   // kernels launched by rtb_render_device calls on this scene
    rtb::WfStreams wf_streams;               // slot-group streams of the wavefront integrator
    bool wf_streams_ok = false;
    std::vector<cudaEvent_t> stage_events;   // RTB_RENDER_STAGE_TIMES: events of the last such frame
    bool stage_events_valid = false;
    ~rtb_scene() {
        cudaSetDevice(device);
        cudaDeviceSynchronize();   // nothing queued may still use the blocks that go back to the cache
        for (auto &a : allocs) rtb::g_blocks.release(device, a.first, a.second);
        rtb::g_blocks.release(device, stage, stage_bytes);
        rtb::g_blocks.release(device, wf_state, wf_bytes);
        for (int k = 0; k < 3; ++k) rtb::g_blocks.release(device, scratch[k], scratch_bytes[k]);
        for (auto &e : stage_events) cudaEventDestroy(e);
        if (wf_streams.graph_exec) cudaGraphExecDestroy(wf_streams.graph_exec);
        if (wf_streams.graph) cudaGraphDestroy(wf_streams.graph);
        if (wf_streams_ok) {
            for (auto &s : wf_streams.stream) cudaStreamDestroy(s);
            for (auto &e : wf_streams.join) cudaEventDestroy(e);
            cudaEventDestroy(wf_streams.fork);
        }
        if (queue_heads) cudaFree(queue_heads);
        for (auto &s : streams) if (s) cudaStreamDestroy(s);
        for (auto &e : ev) if (e) cudaEventDestroy(e);
    }
};

namespace {

using namespace rtb;

// The whole scene goes into ONE device allocation (256-byte aligned pieces): one cudaMalloc / cudaFree instead of ~80.
// The copy is
// staged through two pinned buffers per host thread (allocated once, reused by every later scene): a few host threads
// fill one buffer while the DMA engine drains the other.  A pageable cudaMemcpy stages through the driver on one
// thread: 66 ms for the 335 MB Sponza stand-in, measured (tools/e2e_probe.py); this path is part of bench.py's e2e step.
class Uploader {
  public:
    // *dst receives base + offset in alloc(); src == nullptr: the piece is only reserved (a kernel fills it)
    void add(const void *src, size_t bytes, const void **dst) {
        pieces_.push_back({src, bytes, total_, dst});
        total_ += (bytes + 255) / 256 * 256;
    }
    // allocates; pointers are valid (but their contents not yet copied) after this
    // temp: the block is not the scene's; the caller hands it back with release_temp() once the kernels that read it are done
    int alloc(rtb_scene *sc, bool temp = false) {
        for (auto &p : pieces_) *p.dst = nullptr;
        if (total_ == 0) return RTB_OK;
        size_t got = 0;
        RTB_CUDA(rtb::g_blocks.acquire(sc->device, total_, &base_, &got));
        if (temp) got_ = got;
        else sc->allocs.emplace_back(base_, got);
        for (auto &p : pieces_) if (p.bytes) *p.dst = (char *) base_ + p.off;
        return RTB_OK;
    }
    int run(rtb_scene *sc) {
        if (total_ == 0) return RTB_OK;
        constexpr size_t kBuf = (size_t) 32 << 20;
        struct Stage {
            void *buf[2] = {nullptr, nullptr};
            cudaEvent_t ev[2] = {nullptr, nullptr};
            cudaStream_t st = nullptr;
            bool ok = false;
        };
        static thread_local Stage stage;   // (lives as long as the thread: pinned memory is expensive to allocate)
        if (!stage.ok) {
            bool ok = cudaHostAlloc(&stage.buf[0], kBuf, cudaHostAllocPortable) == cudaSuccess &&
                      cudaHostAlloc(&stage.buf[1], kBuf, cudaHostAllocPortable) == cudaSuccess;
            if (!ok) {   // no pinned memory to be had: plain copies
                cudaGetLastError();
                if (stage.buf[0]) { cudaFreeHost(stage.buf[0]); stage.buf[0] = nullptr; }
                stage.buf[1] = nullptr;
                for (auto &p : pieces_) if (p.bytes && p.src) RTB_CUDA(cudaMemcpy((char *) base_ + p.off, p.src, p.bytes, cudaMemcpyHostToDevice));
                for (auto &p : pieces_) if (p.src) sc->upload_bytes += p.bytes;
                return RTB_OK;
            }
            stage.ok = true;
        }
        // streams and events belong to a device: made per call (cheap)
        cudaError_t err = cudaStreamCreateWithFlags(&stage.st, cudaStreamNonBlocking);
        if (err != cudaSuccess) { stage.st = nullptr; RTB_CUDA(err); }
        stage.ev[0] = stage.ev[1] = nullptr;
        for (auto &e : stage.ev) if (err == cudaSuccess) err = cudaEventCreateWithFlags(&e, cudaEventDisableTiming);
        int k = 0;
        bool used[2] = {false, false};
        for (auto &p : pieces_) {
            if (!p.src) continue;
            for (size_t done = 0; done < p.bytes && err == cudaSuccess; done += kBuf) {   // (nothing runs once err is set)
                const size_t m = std::min(kBuf, p.bytes - done);
                if (used[k]) err = cudaEventSynchronize(stage.ev[k]);
                parallel_copy(stage.buf[k], (const char *) p.src + done, m);
                if (err == cudaSuccess) err = cudaMemcpyAsync((char *) base_ + p.off + done, stage.buf[k], m, cudaMemcpyHostToDevice, stage.st);
                if (err == cudaSuccess) err = cudaEventRecord(stage.ev[k], stage.st);
                used[k] = true;
                k ^= 1;
            }
            sc->upload_bytes += p.bytes;
        }
        if (err == cudaSuccess) err = cudaStreamSynchronize(stage.st);
        for (auto &e : stage.ev) if (e) cudaEventDestroy(e);
        cudaStreamDestroy(stage.st);
        RTB_CUDA(err);
        return RTB_OK;
    }

    void release_temp(rtb_scene *sc) {
        if (base_ && got_) rtb::g_blocks.release(sc->device, base_, got_);
        base_ = nullptr;
        got_ = 0;
    }

  private:
    static void parallel_copy(void *dst, const void *src, size_t bytes) {
        constexpr size_t kMinPerThread = (size_t) 2 << 20;
        const unsigned hw = std::max(1u, std::thread::hardware_concurrency());
        const size_t nt = std::min<size_t>(std::min<size_t>(4, hw), std::max<size_t>(1, bytes / kMinPerThread));
        if (nt <= 1) { std::memcpy(dst, src, bytes); return; }
        std::vector<std::thread> th;
        const size_t per = (bytes / nt + 63) / 64 * 64;
        for (size_t t = 0; t < nt; ++t) {
            const size_t b = t * per, e = std::min(bytes, b + per);
            if (b >= e) break;
            th.emplace_back([=] { std::memcpy((char *) dst + b, (const char *) src + b, e - b); });
        }
        for (auto &t : th) t.join();
    }
    struct P {
        const void *src;
        size_t bytes, off;
        const void **dst;
    };
    std::vector<P> pieces_;
    size_t total_ = 0;
    void *base_ = nullptr;
    size_t got_ = 0;   // temp blocks only
};

int ensure_stage(rtb_scene *sc, size_t bytes) {
    if (sc->stage_bytes >= bytes) return RTB_OK;
    if (sc->stage) { RTB_CUDA(cudaDeviceSynchronize()); rtb::g_blocks.release(sc->device, sc->stage, sc->stage_bytes); }
    sc->stage = nullptr;
    sc->stage_bytes = 0;
    size_t got = 0;
    RTB_CUDA(rtb::g_blocks.acquire(sc->device, bytes, &sc->stage, &got));
    sc->stage_bytes = got;
    return RTB_OK;
}

}  // namespace

namespace rtb {
// see BlockCache::release (rtb_dist.cu: the threads of rtb_render_multi)
void blocks_defer_frees(bool on) { BlockCache::defer_flag() = on; }
void blocks_flush_deferred() { g_blocks.flush_deferred(); }
// device scratch block `which` (0..2) of a scene, at least `bytes` long (rtb_dist.cu)
int scene_scratch(rtb_scene *sc, int which, size_t bytes, void **out) {
    if (sc->scratch_bytes[which] < bytes) {
        if (sc->scratch[which]) { RTB_CUDA(cudaDeviceSynchronize()); g_blocks.release(sc->device, sc->scratch[which], sc->scratch_bytes[which]); }
        sc->scratch[which] = nullptr;
        sc->scratch_bytes[which] = 0;
        size_t got = 0;
        RTB_CUDA(g_blocks.acquire(sc->device, bytes, &sc->scratch[which], &got));
        sc->scratch_bytes[which] = got;
    }
    *out = sc->scratch[which];
    return RTB_OK;
}
cudaStream_t scene_stream(rtb_scene *sc) { return sc->streams[0]; }
}  // namespace rtb

namespace {
// queue cursors, streams, events of a new scene
int finish_scene(rtb_scene *sc, const char *who) {
    cudaError_t ce = cudaMalloc((void **) &sc->queue_heads, kQueueHeads * sizeof(unsigned));
    for (auto &s : sc->streams) if (ce == cudaSuccess) ce = cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking);
    for (auto &e : sc->ev) if (ce == cudaSuccess) ce = cudaEventCreate(&e);
    if (ce != cudaSuccess) {
        rtb::set_error(std::string(who) + ": " + cudaGetErrorString(ce));
        return RTB_ERR_CUDA;
    }
    return RTB_OK;
}
}  // namespace

extern "C" {

int rtb_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) {
        cudaGetLastError();
        return 0;
    }
    return n;
}

int rtb_scene_create(const rtb_scene_desc *desc, int device, rtb_scene **out) {
    if (!desc || !out) { rtb::set_error("rtb_scene_create: null argument"); return RTB_ERR_INVALID; }
    *out = nullptr;
    const int ndev = rtb_device_count();
    if (ndev <= 0) { rtb::set_error("no CUDA device available (this library has no CPU fallback)"); return RTB_ERR_NO_DEVICE; }
    if (device < 0 || device >= ndev) { rtb::set_error("rtb_scene_create: device index out of range"); return RTB_ERR_INVALID; }
    RTB_CUDA(cudaSetDevice(device));

    rtb_scene *sc = new rtb_scene();
    sc->device = device;
    std::memset(&sc->d, 0, sizeof(sc->d));
    auto fail = [&](int rc) { delete sc; return rc; };

    std::string err;
    HostLayout L;
    if (!build_layout(desc, L, err)) {
        rtb::set_error("rtb_scene_create: " + err);
        return fail(RTB_ERR_INVALID);
    }
    int rc;
    fill_lut(L.lut);
    std::vector<DTexture> tex(desc->n_textures);
    std::vector<const void *> tex_ptr(desc->n_textures, nullptr), quad_ptr(desc->n_textures, nullptr), fused_ptr(L.fused.size(), nullptr);
    std::vector<DFused> fused(L.fused.size());
    Uploader up;
    auto add = [&up](const auto &v, auto **dst) { up.add(v.data(), v.size() * sizeof(v[0]), (const void **) dst); };
    add(L.bvh.nodes, &sc->d.bvh.nodes);
    add(L.bvh.tris, &sc->d.bvh.tris);
    add(L.bvh.big, &sc->d.bvh.big_leaves);
    add(L.shade, &sc->d.tri_shade);
    add(L.tan, &sc->d.tri_tan);
    add(L.material, &sc->d.mesh_material);
    add(L.normal_transform, &sc->d.mesh_normal_transform);
    if (desc->n_lights > 0) {
        add(L.light_bvh.nodes, &sc->d.light_bvh.nodes);
        add(L.light_bvh.tris, &sc->d.light_bvh.tris);
        add(L.light_bvh.big, &sc->d.light_bvh.big_leaves);
        add(L.light_bvh.inv_area, &sc->d.light_inv_area);
    }
    add(L.lut, &sc->d.lut);
    add(tex, &sc->d.textures);   // (filled in once the texels' device addresses are known, before the copy runs)
    add(fused, &sc->d.fused);
    for (size_t g = 0; g < L.fused.size(); ++g) up.add(nullptr, (size_t) L.fused[g].width * L.fused[g].height * 64, &fused_ptr[g]);
    for (uint32_t i = 0; i < desc->n_textures; ++i) {
        const rtb_texture_desc &t = desc->textures[i];
        up.add(t.rgba, (size_t) t.width * t.height * 4, &tex_ptr[i]);
        if (L.texture_needs_quads[i]) up.add(nullptr, (size_t) t.width * t.height * 16, &quad_ptr[i]);   // DTexture::quads, written on the device
    }
    if ((rc = up.alloc(sc)) != RTB_OK) return fail(rc);
    for (uint32_t i = 0; i < desc->n_textures; ++i) tex[i] = {(const uchar4 *) tex_ptr[i], (const uint4 *) quad_ptr[i], desc->textures[i].width, desc->textures[i].height};
    for (size_t g = 0; g < L.fused.size(); ++g) fused[g] = {(const uint4 *) fused_ptr[g], L.fused[g].width, L.fused[g].height};
    if ((rc = up.run(sc)) != RTB_OK) return fail(rc);
    fill_bvh_scalars(L.bvh, sc->d.bvh, desc->n_tris);
    if (desc->n_lights > 0) fill_bvh_scalars(L.light_bvh, sc->d.light_bvh, desc->n_lights);
    sc->traversal_bytes = L.bvh.nodes.size() * sizeof(DNode) + L.bvh.tris.size() * sizeof(float4);
    sc->n_big = L.bvh.big.size();
    sc->d.n_meshes = L.n_meshes;
    fill_scalars(desc, sc->d);

    // (failures from here on still go through fail(): the scene and its device arena are released)
    if ((rc = finish_scene(sc, "rtb_scene_create")) != RTB_OK) return fail(rc);
    for (const DTexture &t : tex) launch_texture_quads(t, sc->streams[0]);
    for (size_t g = 0; g < L.fused.size(); ++g) {
        auto px = [&](int ti) { return ti >= 0 ? tex[(size_t) ti].texels : nullptr; };
        launch_fuse_textures(fused[g], px(L.fused[g].tex[0]), px(L.fused[g].tex[1]), px(L.fused[g].tex[2]), sc->streams[0]);
    }
    {
        cudaError_t ce = cudaStreamSynchronize(sc->streams[0]);
        if (ce == cudaSuccess) ce = cudaGetLastError();
        if (ce != cudaSuccess) { rtb::set_error(std::string("rtb_scene_create: ") + cudaGetErrorString(ce)); return fail(RTB_ERR_CUDA); }
    }
    *out = sc;
    return RTB_OK;
}

// Device-side assembly (include/rtb.h, csrc/rtb_assemble.cuh): the raw views and the reference-order nodes are uploaded as
// they are into a temporary block; two kernels write the scene's own buffers from them.
int rtb_scene_create_raw(const rtb_scene_desc *desc, const rtb_raw_geometry *raw, int device, rtb_scene **out) {
    if (!desc || !raw || !out) { rtb::set_error("rtb_scene_create_raw: null argument"); return RTB_ERR_INVALID; }
    *out = nullptr;
    const int ndev = rtb_device_count();
    if (ndev <= 0) { rtb::set_error("no CUDA device available (this library has no CPU fallback)"); return RTB_ERR_NO_DEVICE; }
    if (device < 0 || device >= ndev) { rtb::set_error("rtb_scene_create_raw: device index out of range"); return RTB_ERR_INVALID; }
    auto invalid = [](const std::string &m) { rtb::set_error("rtb_scene_create_raw: " + m); return RTB_ERR_INVALID; };
    if (!raw->meshes || raw->n_meshes == 0 || !raw->tri_source) return invalid("no raw meshes / tri_source");
    if (desc->n_tris == 0 || desc->n_tris >= (1u << 28)) return invalid(desc->n_tris ? "too many triangles for the 28-bit leaf link" : "scene without nodes / triangles");

    std::string err;
    HostLayout L;
    if (!build_layout_common(desc, L, err)) return invalid(err);
    // ---- the views: validated, then merged into the byte ranges that are uploaded (interleaved views share their bytes)
    struct Range { const uint8_t *b, *e; };
    std::vector<Range> ranges;
    std::vector<uint32_t> mesh_first(raw->n_meshes + 1, 0);
    for (uint32_t m = 0; m < raw->n_meshes; ++m) {
        const rtb_raw_mesh &r = raw->meshes[m];
        if ((uint64_t) mesh_first[m] + r.n_tris >= (1ull << 32)) return invalid("too many raw triangles");
        mesh_first[m + 1] = mesh_first[m] + r.n_tris;
        if (r.n_tris == 0) continue;
        if (r.mesh_id < 0 || r.mesh_id >= L.n_meshes) return invalid("mesh id out of range");
        if (!r.position || r.n_vertices == 0) return invalid("raw mesh without positions");
        if (r.normal_mode < RTB_NORMALS_TRANSFORM || r.normal_mode > RTB_NORMALS_FLAT || (r.normal_mode == RTB_NORMALS_VERBATIM && !r.normal))
            return invalid("bad normal_mode");
        auto view = [&](const void *p, int stride, size_t elem, size_t count) -> bool {
            if (!p) return true;
            if (stride < (int) elem || stride % 4 != 0 || ((uintptr_t) p & 3u)) return false;
            ranges.push_back({(const uint8_t *) p, (const uint8_t *) p + (count - 1) * (size_t) stride + elem});
            return true;
        };
        if (!view(r.position, r.position_stride, 12, r.n_vertices) || !view(r.normal, r.normal_stride, 12, r.n_vertices) ||
            !view(r.tangent, r.tangent_stride, 16, r.n_vertices) || !view(r.texcoord, r.texcoord_stride, 8, r.n_vertices))
            return invalid("vertex views must be 4-byte aligned with strides that are multiples of 4 (use rtb_scene_create for other files)");
        if (r.indices) {
            if ((r.index_bytes != 2 && r.index_bytes != 4) || r.index_stride < r.index_bytes || r.index_stride % r.index_bytes != 0 ||
                ((uintptr_t) r.indices % (uintptr_t) r.index_bytes))
                return invalid("index view: 2- or 4-byte indices, aligned");
            ranges.push_back({(const uint8_t *) r.indices, (const uint8_t *) r.indices + ((size_t) r.n_tris * 3 - 1) * (size_t) r.index_stride + (size_t) r.index_bytes});
        } else if ((uint64_t) r.n_tris * 3 > r.n_vertices) {
            return invalid("[gltf check] vertex index out of range");
        }
    }
    for (auto &g : ranges) g.b = (const uint8_t *) ((uintptr_t) g.b & ~(uintptr_t) 15);   // (stays inside the page of the first byte)
    std::sort(ranges.begin(), ranges.end(), [](const Range &a, const Range &b) { return a.b < b.b; });
    std::vector<Range> merged;
    for (const Range &g : ranges) {
        if (!merged.empty() && g.b <= merged.back().e + 64) merged.back().e = std::max(merged.back().e, g.e);
        else merged.push_back(g);
    }
    // ---- nodes
    const int32_t *root = desc->node_links;
    const bool root_leaf = root[4] > 0;
    if (!root_leaf && (desc->n_nodes & 1u) == 0) return invalid("nodes are not in the reference's order");
    const uint32_t n_internal = root_leaf ? 0u : (desc->n_nodes - 1u) / 2u;
    LeafLinker leaf;
    big_leaf_slots(desc->n_nodes, desc->node_links, desc->n_tris, L.bvh.big, leaf.big_node, leaf.big_slot);
    if (root_leaf && (root[3] < 0 || (uint64_t) root[3] + (uint64_t) root[4] > desc->n_tris)) return invalid("leaf range out of bounds");

    RTB_CUDA(cudaSetDevice(device));
    rtb_scene *sc = new rtb_scene();
    sc->device = device;
    std::memset(&sc->d, 0, sizeof(sc->d));
    Uploader up, tmp;
    auto fail = [&](int rc) { tmp.release_temp(sc); delete sc; return rc; };
    int rc;
    fill_lut(L.lut);
    std::vector<DTexture> tex(desc->n_textures);
    std::vector<const void *> tex_ptr(desc->n_textures, nullptr), quad_ptr(desc->n_textures, nullptr), fused_ptr(L.fused.size(), nullptr);
    std::vector<DFused> fused(L.fused.size());
    auto add = [&up](const auto &v, auto **dst) { up.add(v.data(), v.size() * sizeof(v[0]), (const void **) dst); };
    const size_t nt = desc->n_tris;
    up.add(nullptr, (size_t) n_internal * sizeof(DNode), (const void **) &sc->d.bvh.nodes);
    up.add(nullptr, nt * kTriStride * sizeof(float4), (const void **) &sc->d.bvh.tris);
    add(L.bvh.big, &sc->d.bvh.big_leaves);
    up.add(nullptr, nt * 4 * sizeof(float4), (const void **) &sc->d.tri_shade);
    if (L.any_normal_map) up.add(nullptr, nt * 3 * sizeof(float4), (const void **) &sc->d.tri_tan);
    add(L.material, &sc->d.mesh_material);
    add(L.normal_transform, &sc->d.mesh_normal_transform);
    if (desc->n_lights > 0) {
        add(L.light_bvh.nodes, &sc->d.light_bvh.nodes);
        add(L.light_bvh.tris, &sc->d.light_bvh.tris);
        add(L.light_bvh.big, &sc->d.light_bvh.big_leaves);
        add(L.light_bvh.inv_area, &sc->d.light_inv_area);
    }
    add(L.lut, &sc->d.lut);
    add(tex, &sc->d.textures);
    add(fused, &sc->d.fused);
    for (size_t g = 0; g < L.fused.size(); ++g) up.add(nullptr, (size_t) L.fused[g].width * L.fused[g].height * 64, &fused_ptr[g]);
    for (uint32_t i = 0; i < desc->n_textures; ++i) {
        const rtb_texture_desc &t = desc->textures[i];
        up.add(t.rgba, (size_t) t.width * t.height * 4, &tex_ptr[i]);
        if (L.texture_needs_quads[i]) up.add(nullptr, (size_t) t.width * t.height * 16, &quad_ptr[i]);   // DTexture::quads, written on the device
    }
    if ((rc = up.alloc(sc)) != RTB_OK) return fail(rc);
    for (uint32_t i = 0; i < desc->n_textures; ++i) tex[i] = {(const uchar4 *) tex_ptr[i], (const uint4 *) quad_ptr[i], desc->textures[i].width, desc->textures[i].height};
    for (size_t g = 0; g < L.fused.size(); ++g) fused[g] = {(const uint4 *) fused_ptr[g], L.fused[g].width, L.fused[g].height};

    // ---- the temporary block: raw bytes, reference-order nodes, tables, the two node frontiers
    std::vector<DRawMesh> dmesh(raw->n_meshes);
    std::vector<const void *> range_dev(merged.size(), nullptr);
    const float *d_aabb = nullptr;
    const int32_t *d_links = nullptr;
    const uint32_t *d_source = nullptr, *d_first = nullptr, *d_big_node = nullptr, *d_big_slot = nullptr;
    const DRawMesh *d_mesh = nullptr;
    const NodeItem *d_items = nullptr;
    const unsigned *d_words = nullptr;   // [0,1]: frontier sizes, [2]: error flag
    for (size_t k = 0; k < merged.size(); ++k) tmp.add(merged[k].b, (size_t) (merged[k].e - merged[k].b), &range_dev[k]);
    tmp.add(raw->tri_source, nt * sizeof(uint32_t), (const void **) &d_source);
    tmp.add(mesh_first.data(), mesh_first.size() * sizeof(uint32_t), (const void **) &d_first);
    tmp.add(dmesh.data(), dmesh.size() * sizeof(DRawMesh), (const void **) &d_mesh);
    if (n_internal > 0) {
        tmp.add(desc->node_aabb, (size_t) desc->n_nodes * 6 * sizeof(float), (const void **) &d_aabb);
        tmp.add(desc->node_links, (size_t) desc->n_nodes * 5 * sizeof(int32_t), (const void **) &d_links);
        tmp.add(leaf.big_node.data(), leaf.big_node.size() * sizeof(uint32_t), (const void **) &d_big_node);
        tmp.add(leaf.big_slot.data(), leaf.big_slot.size() * sizeof(uint32_t), (const void **) &d_big_slot);
        tmp.add(nullptr, (size_t) n_internal * 2 * sizeof(NodeItem), (const void **) &d_items);
    }
    tmp.add(nullptr, 256, (const void **) &d_words);
    if ((rc = tmp.alloc(sc, true)) != RTB_OK) return fail(rc);
    auto to_dev = [&](const void *p) -> const uint8_t * {
        if (!p) return nullptr;
        // the merged range that holds p: the last one starting at or before it
        size_t lo = 0, hi = merged.size();
        while (hi - lo > 1) {
            const size_t mid = (lo + hi) / 2;
            if (merged[mid].b <= (const uint8_t *) p) lo = mid; else hi = mid;
        }
        return (const uint8_t *) range_dev[lo] + ((const uint8_t *) p - merged[lo].b);
    };
    for (uint32_t m = 0; m < raw->n_meshes; ++m) {
        const rtb_raw_mesh &r = raw->meshes[m];
        DRawMesh &d = dmesh[m];
        std::memset(&d, 0, sizeof(d));
        if (r.n_tris == 0) continue;
        d.idx = to_dev(r.indices); d.pos = to_dev(r.position); d.nrm = to_dev(r.normal); d.tan = to_dev(r.tangent); d.uv = to_dev(r.texcoord);
        d.first_tri = mesh_first[m]; d.n_tris = r.n_tris; d.n_vertices = r.n_vertices;
        d.index_bytes = r.index_bytes; d.idx_stride = r.index_stride;
        d.pos_stride = r.position_stride; d.nrm_stride = r.normal_stride; d.tan_stride = r.tangent_stride; d.uv_stride = r.texcoord_stride;
        d.normal_mode = r.normal_mode; d.flags = r.flags; d.mesh_id = r.mesh_id;
        std::memcpy(d.transform, r.transform, sizeof(d.transform));
        std::memcpy(d.normal_transform, &L.normal_transform[(size_t) r.mesh_id * 16], sizeof(d.normal_transform));
    }
    // (failures from here on still go through fail(): the scene, its arena and the temporary block are released)
    if ((rc = finish_scene(sc, "rtb_scene_create_raw")) != RTB_OK) return fail(rc);
    if ((rc = up.run(sc)) != RTB_OK) return fail(rc);
    if ((rc = tmp.run(sc)) != RTB_OK) return fail(rc);
    cudaStream_t st = sc->streams[0];
    unsigned *words = const_cast<unsigned *>(d_words);
    int *d_error = (int *) (words + 2);
    cudaError_t ce = cudaMemsetAsync(words, 0, 256, st);
    int levels = 0;
    if (ce == cudaSuccess) {
        for (const DTexture &t : tex) launch_texture_quads(t, st);
        for (size_t g = 0; g < L.fused.size(); ++g) {
            auto px = [&](int ti) { return ti >= 0 ? tex[(size_t) ti].texels : nullptr; };
            launch_fuse_textures(fused[g], px(L.fused[g].tex[0]), px(L.fused[g].tex[1]), px(L.fused[g].tex[2]), st);
        }
        launch_assemble_triangles(d_mesh, d_first, (int) raw->n_meshes, d_source, desc->n_tris, const_cast<float4 *>(sc->d.bvh.tris),
                                  const_cast<float4 *>(sc->d.tri_shade), const_cast<float4 *>(sc->d.tri_tan), d_error, st);
        const BigLeafIndex big{d_big_node, d_big_slot, (int) leaf.big_node.size()};
        levels = launch_convert_nodes(desc->n_nodes, n_internal, desc->n_tris, d_aabb, d_links, big, const_cast<DNode *>(sc->d.bvh.nodes),
                                      const_cast<NodeItem *>(d_items), words, d_error, st);
        sc->launches = 0;
    }
    int flagged = 0;
    if (ce == cudaSuccess) ce = cudaMemcpyAsync(&flagged, d_error, sizeof(int), cudaMemcpyDeviceToHost, st);
    if (ce == cudaSuccess) ce = cudaStreamSynchronize(st);
    if (ce == cudaSuccess) ce = cudaGetLastError();
    if (ce != cudaSuccess || levels < 0) {
        rtb::set_error(std::string("rtb_scene_create_raw: ") + (ce != cudaSuccess ? cudaGetErrorString(ce) : "node conversion failed"));
        return fail(RTB_ERR_CUDA);
    }
    if (flagged != kAsmOk) { rtb::set_error(std::string("rtb_scene_create_raw: ") + assemble_error_text(flagged)); return fail(RTB_ERR_INVALID); }
    if (levels > kMaxStack) {
        rtb::set_error("rtb_scene_create_raw: BVH deeper than the traversal stack (" + std::to_string(levels) + " > " + std::to_string(kMaxStack) + ")");
        return fail(RTB_ERR_INVALID);
    }
    tmp.release_temp(sc);
    std::memcpy(sc->d.bvh.root_box, desc->node_aabb, 24);
    sc->d.bvh.root_link = root_leaf ? leaf.link(0, root[3], root[4]) : 0;
    sc->d.bvh.n_internal = (int) n_internal;
    sc->d.bvh.n_tris = (int) desc->n_tris;
    sc->d.bvh.max_depth = levels;
    if (desc->n_lights > 0) fill_bvh_scalars(L.light_bvh, sc->d.light_bvh, desc->n_lights);
    sc->traversal_bytes = (size_t) n_internal * sizeof(DNode) + nt * kTriStride * sizeof(float4);
    sc->n_big = L.bvh.big.size();
    sc->d.n_meshes = L.n_meshes;
    fill_scalars(desc, sc->d);
    *out = sc;
    return RTB_OK;
}

int rtb_scene_read_buffer(rtb_scene *scene, int which, void *dst, size_t *bytes) {
    if (!scene || !bytes) { rtb::set_error("rtb_scene_read_buffer: null argument"); return RTB_ERR_INVALID; }
    RTB_CUDA(cudaSetDevice(scene->device));
    const DBvh &b = scene->d.bvh;
    const void *src = nullptr;
    size_t size = 0;
    struct { float box[6]; int32_t v[4]; } scalars;
    switch (which) {
        case RTB_BUF_NODES: src = b.nodes; size = (size_t) b.n_internal * sizeof(DNode); break;
        case RTB_BUF_TRIS: src = b.tris; size = (size_t) b.n_tris * kTriStride * sizeof(float4); break;
        case RTB_BUF_SHADE: src = scene->d.tri_shade; size = (size_t) b.n_tris * 4 * sizeof(float4); break;
        case RTB_BUF_TANGENTS: src = scene->d.tri_tan; size = scene->d.tri_tan ? (size_t) b.n_tris * 3 * sizeof(float4) : 0; break;
        case RTB_BUF_BIG_LEAVES: src = b.big_leaves; size = scene->n_big * sizeof(int2); break;
        case RTB_BUF_BVH_SCALARS:
            std::memcpy(scalars.box, b.root_box, 24);
            scalars.v[0] = b.root_link; scalars.v[1] = b.n_internal; scalars.v[2] = b.n_tris; scalars.v[3] = b.max_depth;
            size = sizeof(scalars);
            if (dst) std::memcpy(dst, &scalars, std::min(*bytes, size));
            *bytes = size;
            return RTB_OK;
        default: rtb::set_error("rtb_scene_read_buffer: unknown buffer"); return RTB_ERR_INVALID;
    }
    if (dst && size && *bytes) RTB_CUDA(cudaMemcpy(dst, src, std::min(*bytes, size), cudaMemcpyDeviceToHost));
    *bytes = size;
    return RTB_OK;
}

int rtb_measure_l2(int device, size_t bytes, double *stream_gbs, double *gather64_gbs) {
    if (rtb_device_count() <= 0) { rtb::set_error("no CUDA device available (this library has no CPU fallback)"); return RTB_ERR_NO_DEVICE; }
    RTB_CUDA(cudaSetDevice(device));
    if (rtb::measure_l2(bytes, stream_gbs, gather64_gbs) != 0) { rtb::set_error("rtb_measure_l2: CUDA error"); return RTB_ERR_CUDA; }
    return RTB_OK;
}

void rtb_scene_destroy(rtb_scene *scene) { delete scene; }
int rtb_scene_device(const rtb_scene *scene) { return scene ? scene->device : -1; }
size_t rtb_scene_traversal_bytes(const rtb_scene *scene) { return scene ? scene->traversal_bytes : 0; }
size_t rtb_scene_upload_bytes(const rtb_scene *scene) { return scene ? scene->upload_bytes : 0; }

int rtb_closest_hit_device(rtb_scene *scene, const float *d_rays6, size_t n, int cull_mode, int32_t *d_id, float *d_t, float *d_u,
                           float *d_v, uint8_t *d_inside, unsigned long long *d_counters, void *stream) {
    if (!scene || !d_rays6 || !d_id) { rtb::set_error("rtb_closest_hit_device: null argument"); return RTB_ERR_INVALID; }
    if (cull_mode != RTB_CULL_REFERENCE && cull_mode != RTB_CULL_GLOBAL) { rtb::set_error("rtb_closest_hit_device: bad cull mode"); return RTB_ERR_INVALID; }
    RTB_CUDA(cudaSetDevice(scene->device));
    rtb::launch_closest_hit(scene->d, d_rays6, n, cull_mode, d_id, d_t, d_u, d_v, d_inside, d_counters,
                            scene->queue_heads + (scene->next_head++ % kQueueHeads), (cudaStream_t) stream);
    RTB_CUDA(cudaGetLastError());
    return RTB_OK;
}

// rtb_closest_hit_device for a caller that must not allocate device memory (include/rtb.h): the queue cursor comes from a
// table compiled into the library (csrc/rtb_kernels.cu), nothing on this path calls cudaMalloc / cudaFree.
int rtb_closest_hit_static(rtb_scene *scene, const float *d_rays6, size_t n, int cull_mode, int32_t *d_id, float *d_t, float *d_u,
                           float *d_v, uint8_t *d_inside, unsigned long long *d_counters, void *stream) {
    if (!scene || (n > 0 && (!d_rays6 || !d_id))) { rtb::set_error("rtb_closest_hit_static: null argument"); return RTB_ERR_INVALID; }
    if (cull_mode != RTB_CULL_REFERENCE && cull_mode != RTB_CULL_GLOBAL) { rtb::set_error("rtb_closest_hit_static: bad cull mode"); return RTB_ERR_INVALID; }
    if (n > (size_t) RTB_CLOSEST_HIT_STATIC_MAX_RAYS) {
        rtb::set_error("rtb_closest_hit_static: " + std::to_string(n) + " rays in one call; the built-in 32-bit ray cursor serves at most " +
                       std::to_string((unsigned long long) RTB_CLOSEST_HIT_STATIC_MAX_RAYS) + " (split the batch)");
        return RTB_ERR_INVALID;
    }
    RTB_CUDA(cudaSetDevice(scene->device));
    const cudaError_t e = rtb::launch_closest_hit_static(scene->d, scene->device, d_rays6, n, cull_mode, d_id, d_t, d_u, d_v, d_inside,
                                                         d_counters, (cudaStream_t) stream);
    if (e != cudaSuccess) {
        rtb::set_error(std::string("rtb_closest_hit_static: ") + cudaGetErrorString(e));
        return RTB_ERR_CUDA;
    }
    return RTB_OK;
}

// Host-pointer variant: the batch is cut into chunks that flow through three streams so that the host->device copy of
// chunk k+1 and the device->host copy of chunk k-1 overlap the traversal of chunk k.
int rtb_closest_hit(rtb_scene *scene, const float *rays6, size_t n, int cull_mode, int32_t *id, float *t, float *u, float *v,
                    uint8_t *inside, rtb_stats *stats) {
    if (!scene || !id || (n > 0 && !rays6)) { rtb::set_error("rtb_closest_hit: null argument"); return RTB_ERR_INVALID; }
    if (cull_mode != RTB_CULL_REFERENCE && cull_mode != RTB_CULL_GLOBAL) { rtb::set_error("rtb_closest_hit: bad cull mode"); return RTB_ERR_INVALID; }
    RTB_CUDA(cudaSetDevice(scene->device));
    const bool count = stats != nullptr && stats->box_tests == RTB_REQUEST_COUNTERS;
    if (stats) std::memset(stats, 0, sizeof(*stats));
    if (n == 0) return RTB_OK;
    const size_t chunk = std::min<size_t>(n, (size_t) 1 << 22);
    const int nslots = 3;
    const size_t per_ray = 24 + 4 + 4 + 4 + 4 + 1;
    const size_t slot_bytes = ((chunk * per_ray + 255) / 256 + 8) * 256;
    int rc = ensure_stage(scene, slot_bytes * nslots + 256);
    if (rc != RTB_OK) return rc;
    unsigned long long *d_cnt = nullptr;
    if (count) {
        d_cnt = (unsigned long long *) ((char *) scene->stage + slot_bytes * nslots);
        RTB_CUDA(cudaMemsetAsync(d_cnt, 0, 16, scene->streams[0]));
        RTB_CUDA(cudaStreamSynchronize(scene->streams[0]));
    }
    RTB_CUDA(cudaEventRecord(scene->ev[0], scene->streams[0]));
    size_t launches = 0;
    for (size_t off = 0, k = 0; off < n; off += chunk, ++k) {
        const size_t m = std::min(chunk, n - off);
        cudaStream_t st = scene->streams[k % nslots];
        char *base = (char *) scene->stage + slot_bytes * (k % nslots);
        float *d_rays = (float *) base;
        int *d_id = (int *) (base + chunk * 24);
        float *d_t = (float *) (base + chunk * 28);
        float *d_u = (float *) (base + chunk * 32);
        float *d_v = (float *) (base + chunk * 36);
        uint8_t *d_in = (uint8_t *) (base + chunk * 40);
        RTB_CUDA(cudaMemcpyAsync(d_rays, rays6 + off * 6, m * 24, cudaMemcpyHostToDevice, st));
        rtb::launch_closest_hit(scene->d, d_rays, m, cull_mode, d_id, t ? d_t : nullptr, u ? d_u : nullptr, v ? d_v : nullptr,
                                inside ? d_in : nullptr, d_cnt, scene->queue_heads + (scene->next_head++ % kQueueHeads), st);
        ++launches;
        RTB_CUDA(cudaMemcpyAsync(id + off, d_id, m * 4, cudaMemcpyDeviceToHost, st));
        if (t) RTB_CUDA(cudaMemcpyAsync(t + off, d_t, m * 4, cudaMemcpyDeviceToHost, st));
        if (u) RTB_CUDA(cudaMemcpyAsync(u + off, d_u, m * 4, cudaMemcpyDeviceToHost, st));
        if (v) RTB_CUDA(cudaMemcpyAsync(v + off, d_v, m * 4, cudaMemcpyDeviceToHost, st));
        if (inside) RTB_CUDA(cudaMemcpyAsync(inside + off, d_in, m, cudaMemcpyDeviceToHost, st));
    }
    for (auto &s : scene->streams) RTB_CUDA(cudaStreamSynchronize(s));
    RTB_CUDA(cudaEventRecord(scene->ev[1], scene->streams[0]));
    RTB_CUDA(cudaEventSynchronize(scene->ev[1]));
    RTB_CUDA(cudaGetLastError());
    if (stats) {
        float ms = 0.f;
        cudaEventElapsedTime(&ms, scene->ev[0], scene->ev[1]);
        stats->kernel_ms = ms;   // whole pipelined call (copies overlapped with traversal)
        stats->closest_hit_rays = n;
        stats->launches = launches;
        if (count) {
            unsigned long long c[2];
            RTB_CUDA(cudaMemcpy(c, d_cnt, 16, cudaMemcpyDeviceToHost));
            stats->box_tests = c[0];
            stats->tri_tests = c[1];
        }
    }
    return RTB_OK;
}

int rtb_light_pdf(rtb_scene *scene, const float *rays6, size_t n, float *pdf) {
    if (!scene || !pdf || (n > 0 && !rays6)) { rtb::set_error("rtb_light_pdf: null argument"); return RTB_ERR_INVALID; }
    RTB_CUDA(cudaSetDevice(scene->device));
    if (n == 0) return RTB_OK;
    int rc = ensure_stage(scene, n * 28 + 256);
    if (rc != RTB_OK) return rc;
    float *d_rays = (float *) scene->stage;
    float *d_pdf = d_rays + n * 6;
    cudaStream_t st = scene->streams[0];
    RTB_CUDA(cudaMemcpyAsync(d_rays, rays6, n * 24, cudaMemcpyHostToDevice, st));
    rtb::launch_light_pdf(scene->d, d_rays, n, d_pdf, st);
    RTB_CUDA(cudaMemcpyAsync(pdf, d_pdf, n * 4, cudaMemcpyDeviceToHost, st));
    RTB_CUDA(cudaStreamSynchronize(st));
    RTB_CUDA(cudaGetLastError());
    return RTB_OK;
}

int rtb_render_local_rows(const rtb_render_params *p) {
    if (!p || p->width <= 0 || p->height <= 0 || p->row_begin < 0 || p->row_end > p->height || p->row_begin > p->row_end) return -1;
    const int rows = p->row_end - p->row_begin;
    if (p->interleave_count <= 1) return rows;
    const int R = p->interleave_rows;
    if (R <= 0 || R % 4 != 0 || p->interleave_index < 0 || p->interleave_index >= p->interleave_count) return -1;
    const int blocks = (rows + R - 1) / R;
    int local = 0;
    for (int b = p->interleave_index; b < blocks; b += p->interleave_count) local += std::min(R, rows - b * R);
    return local;
}

int rtb_render_global_row(const rtb_render_params *p, int lr) {
    if (!p || lr < 0) return -1;
    if (p->interleave_count <= 1) return p->row_begin + lr;
    const int R = p->interleave_rows;
    return p->row_begin + ((lr / R) * p->interleave_count + p->interleave_index) * R + lr % R;
}

int rtb_render_resume_device(rtb_scene *scene, const rtb_render_params *p, void *d_pixel_state, float *d_accum, uint8_t *d_rgb8,
                             unsigned long long *d_counters, void *stream) {
    if (!scene || !p) { rtb::set_error("rtb_render_device: null argument"); return RTB_ERR_INVALID; }
    const int local_rows = rtb_render_local_rows(p);
    if (local_rows < 0 || p->samples <= 0) {
        rtb::set_error("rtb_render_device: bad canvas / row range / interleave / sample count");
        return RTB_ERR_INVALID;
    }
    if (p->sample_begin < 0 || p->sample_begin >= p->samples || (p->sample_begin > 0 && !d_pixel_state)) {
        rtb::set_error("rtb_render_device: sample_begin must be 0, or the sample count of the pixel state handed to rtb_render_resume (< samples): "
                       "a pixel's samples share one sequential stream");
        return RTB_ERR_INVALID;
    }
    RTB_CUDA(cudaSetDevice(scene->device));
    rtb::RenderLaunch L;
    L.width = p->width; L.height = p->height; L.samples = p->samples;
    L.row_begin = p->row_begin; L.row_end = p->row_end;
    L.il_count = p->interleave_count; L.il_index = p->interleave_index; L.il_rows = p->interleave_rows > 0 ? p->interleave_rows : 4;
    L.local_rows = local_rows;
    L.pixel0_seed = p->pixel0_seed;
    L.d_accum = d_accum; L.d_rgb8 = d_rgb8; L.d_counters = d_counters;
    L.count_tests = (p->flags & RTB_RENDER_COUNTERS) != 0 && d_counters != nullptr;
    L.cull_mode = (p->flags & RTB_RENDER_REFERENCE_CULL) ? RTB_CULL_REFERENCE : RTB_CULL_GLOBAL;
    L.sample_begin = p->sample_begin;
    L.d_pixel_state = (uint4 *) d_pixel_state;
    L.substreams = std::max(1, p->substreams);
    L.sub_begin = 0;
    L.sub_end = L.substreams;
    L.plane_slots = 0;
    if (L.substreams > 1) {
        if (p->substream_begin != 0 || p->substream_end != 0) { L.sub_begin = p->substream_begin; L.sub_end = p->substream_end; }
        if (L.sub_begin < 0 || L.sub_begin >= L.sub_end || L.sub_end > L.substreams) {
            rtb::set_error("rtb_render_device: substream range outside [0, substreams)");
            return RTB_ERR_INVALID;
        }
        if (d_pixel_state || p->sample_begin > 0 || (p->flags & (RTB_RENDER_REFERENCE_CULL | RTB_RENDER_MEGAKERNEL))) {
            rtb::set_error("rtb_render_device: substreams > 1 runs on the wavefront integrator only and cannot be resumed");
            return RTB_ERR_INVALID;
        }
        if (d_rgb8 && L.sub_end - L.sub_begin != L.substreams) {
            rtb::set_error("rtb_render_device: a call that holds only part of a pixel's streams returns partial sums; rgb8 must be NULL");
            return RTB_ERR_INVALID;
        }
        if ((unsigned long long) L.substreams * (unsigned long long) p->width * (unsigned long long) p->height >= 2147483646ull) {
            rtb::set_error("rtb_render_device: substreams x width x height must stay below the engine's modulus");
            return RTB_ERR_INVALID;
        }
    }
    // Integrator: the wavefront (ray-queue) form unless the caller asks for the megakernel; RTB_CULL_REFERENCE only
    // exists as a megakernel (the queue traverses with global-best culling).
    const bool mega = (p->flags & (RTB_RENDER_REFERENCE_CULL | RTB_RENDER_MEGAKERNEL)) != 0 || scene->d.ray_depth < 1 || scene->d.ray_depth > 15;
    if (mega && d_pixel_state) {
        rtb::set_error("rtb_render_resume: the pixel state is kept by the wavefront integrator only (no RTB_RENDER_MEGAKERNEL / RTB_RENDER_REFERENCE_CULL)");
        return RTB_ERR_INVALID;
    }
    if (mega && L.substreams > 1) {
        rtb::set_error("rtb_render_device: substreams > 1 needs ray_depth in [1, 15] (wavefront integrator)");
        return RTB_ERR_INVALID;
    }
    if (mega || local_rows == 0) {
        rtb::launch_render(scene->d, L, (cudaStream_t) stream);
        scene->launches += local_rows ? 1 : 0;
    } else {
        const size_t need = rtb::wf_alloc_bytes(rtb::wf_slots(L), std::max(1, scene->d.ray_depth - 1), scene->d.bvh.max_depth);
        if (scene->wf_bytes < need) {
            // the previous state may still be in use by work queued on a caller's stream
            if (scene->wf_state) { RTB_CUDA(cudaDeviceSynchronize()); rtb::g_blocks.release(scene->device, scene->wf_state, scene->wf_bytes); }
            scene->wf_state = nullptr;
            scene->wf_bytes = 0;
            size_t got = 0;
            RTB_CUDA(rtb::g_blocks.acquire(scene->device, need, &scene->wf_state, &got));
            scene->wf_bytes = got;
        }
        const bool timed = (p->flags & RTB_RENDER_STAGE_TIMES) != 0;
        if (!scene->wf_streams_ok) {
            for (auto &s : scene->wf_streams.stream) RTB_CUDA(cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking));
            for (auto &e : scene->wf_streams.join) RTB_CUDA(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
            RTB_CUDA(cudaEventCreateWithFlags(&scene->wf_streams.fork, cudaEventDisableTiming));
            scene->wf_streams_ok = true;
        }
        const int wrc = rtb::launch_render_wavefront(scene->d, L, scene->wf_state, (cudaStream_t) stream, &scene->launches,
                                                     timed ? &scene->stage_events : nullptr, &scene->wf_streams);
        scene->stage_events_valid = timed && wrc == 0;
        if (wrc != 0) {
            const cudaError_t e = cudaGetLastError();
            rtb::set_error(wrc == 2 ? "rtb_render_device: the frame did not finish (slots still live after the bounded number of extra rounds)"
                                    : std::string("rtb_render_device: stream synchronisation failed: ") + cudaGetErrorString(e));
            return RTB_ERR_CUDA;
        }
    }
    RTB_CUDA(cudaGetLastError());
    return RTB_OK;
}

int rtb_render_device(rtb_scene *scene, const rtb_render_params *p, float *d_accum, uint8_t *d_rgb8, unsigned long long *d_counters,
                      void *stream) {
    return rtb_render_resume_device(scene, p, nullptr, d_accum, d_rgb8, d_counters, stream);
}

int rtb_scene_set_progress(rtb_scene *scene, rtb_progress_fn fn, void *user) {
    if (!scene) { rtb::set_error("rtb_scene_set_progress: null scene"); return RTB_ERR_INVALID; }
    scene->wf_streams.progress = fn;
    scene->wf_streams.progress_user = user;
    return RTB_OK;
}

size_t rtb_render_state_bytes(const rtb_render_params *p) {
    const int rows = p ? rtb_render_local_rows(p) : -1;
    return rows < 0 ? 0 : (size_t) rows * (size_t) p->width * RTB_PIXEL_STATE_BYTES;
}

int rtb_render(rtb_scene *scene, const rtb_render_params *p, float *accum, uint8_t *rgb8, rtb_stats *stats) {
    return rtb_render_resume(scene, p, nullptr, accum, rgb8, stats);
}

int rtb_render_resume(rtb_scene *scene, const rtb_render_params *p, void *state, float *accum, uint8_t *rgb8, rtb_stats *stats) {
    if (!scene || !p) { rtb::set_error("rtb_render: null argument"); return RTB_ERR_INVALID; }
    if (stats) std::memset(stats, 0, sizeof(*stats));
    RTB_CUDA(cudaSetDevice(scene->device));
    const int local_rows = rtb_render_local_rows(p);
    if (local_rows < 0) { rtb::set_error("rtb_render: bad canvas / row range / interleave"); return RTB_ERR_INVALID; }
    const size_t npix = (size_t) local_rows * (size_t) p->width;
    const size_t accum_bytes = npix * 3 * sizeof(float), rgb_bytes = npix * 3;
    const size_t state_bytes = state ? npix * RTB_PIXEL_STATE_BYTES : 0;
    const size_t rgb_off = (accum_bytes + 255) / 256 * 256, cnt_off = rgb_off + (rgb_bytes + 255) / 256 * 256;
    const size_t state_off = cnt_off + 256;
    if (state && p->sample_begin > 0 && npix > 0 && ((const uint32_t *) state)[3] != (uint32_t) p->sample_begin) {
        rtb::set_error("rtb_render_resume: the pixel state holds " + std::to_string(((const uint32_t *) state)[3]) +
                       " samples per pixel, sample_begin says " + std::to_string(p->sample_begin));
        return RTB_ERR_INVALID;
    }
    int rc = ensure_stage(scene, state_off + state_bytes);
    if (rc != RTB_OK) return rc;
    char *base = (char *) scene->stage;
    float *d_accum = (float *) base;
    uint8_t *d_rgb = (uint8_t *) (base + rgb_off);
    unsigned long long *d_cnt = (unsigned long long *) (base + cnt_off);
    cudaStream_t st = scene->streams[0];
    RTB_CUDA(cudaMemsetAsync(d_cnt, 0, 48, st));
    RTB_CUDA(cudaEventRecord(scene->ev[0], st));
    const unsigned long long launches_before = scene->launches;
    void *d_state = state ? (void *) (base + state_off) : nullptr;
    if (state && p->sample_begin > 0) RTB_CUDA(cudaMemcpyAsync(d_state, state, state_bytes, cudaMemcpyHostToDevice, st));
    rc = rtb_render_resume_device(scene, p, d_state, d_accum, rgb8 ? d_rgb : nullptr, d_cnt, st);
    if (rc != RTB_OK) return rc;
    RTB_CUDA(cudaEventRecord(scene->ev[1], st));
    if (state && state_bytes) RTB_CUDA(cudaMemcpyAsync(state, d_state, state_bytes, cudaMemcpyDeviceToHost, st));
    if (accum && accum_bytes) RTB_CUDA(cudaMemcpyAsync(accum, d_accum, accum_bytes, cudaMemcpyDeviceToHost, st));
    if (rgb8 && rgb_bytes) RTB_CUDA(cudaMemcpyAsync(rgb8, d_rgb, rgb_bytes, cudaMemcpyDeviceToHost, st));
    unsigned long long c[6] = {0, 0, 0, 0, 0, 0};
    RTB_CUDA(cudaMemcpyAsync(c, d_cnt, 48, cudaMemcpyDeviceToHost, st));
    RTB_CUDA(cudaStreamSynchronize(st));
    RTB_CUDA(cudaGetLastError());
    if (stats) {
        float ms = 0.f;
        cudaEventElapsedTime(&ms, scene->ev[0], scene->ev[1]);
        stats->kernel_ms = ms;
        stats->closest_hit_rays = c[0];
        stats->light_pdf_queries = c[1];
        stats->box_tests = c[2];
        stats->tri_tests = c[3];
        stats->light_box_tests = c[4];
        stats->light_tri_tests = c[5];
        stats->launches = scene->launches - launches_before;
    }
    return RTB_OK;
}

void rtb_release_cached_memory(void) { rtb::g_blocks.flush_deferred(); rtb::g_blocks.trim(-1); }

unsigned long long rtb_scene_launches(const rtb_scene *scene) { return scene ? scene->launches : 0ull; }

int rtb_scene_stage_ms(rtb_scene *scene, double *trace_ms, double *other_ms, unsigned long long *trace_launches) {
    if (!scene || !scene->stage_events_valid || scene->stage_events.size() < 3) {
        rtb::set_error("rtb_scene_stage_ms: the last render on this scene did not ask for RTB_RENDER_STAGE_TIMES");
        return RTB_ERR_INVALID;
    }
    RTB_CUDA(cudaSetDevice(scene->device));
    const auto &ev = scene->stage_events;
    RTB_CUDA(cudaEventSynchronize(ev.back()));
    double tr = 0.0, ot = 0.0;
    const size_t iters = (ev.size() - 1) / 2;
    for (size_t i = 0; i < iters; ++i) {
        float a = 0.f, b = 0.f;
        RTB_CUDA(cudaEventElapsedTime(&a, ev[2 * i], ev[2 * i + 1]));
        RTB_CUDA(cudaEventElapsedTime(&b, ev[2 * i + 1], ev[2 * i + 2]));
        tr += a;
        ot += b;
    }
    if (trace_ms) *trace_ms = tr;
    if (other_ms) *other_ms = ot;
    if (trace_launches) *trace_launches = iters;
    return RTB_OK;
}

int rtb_resolve_device(const float *d_accum, size_t n_pixels, int samples, uint8_t *d_rgb8, void *stream) {
    if (!d_accum || !d_rgb8 || samples <= 0) { rtb::set_error("rtb_resolve_device: invalid argument"); return RTB_ERR_INVALID; }
    rtb::launch_resolve(d_accum, n_pixels, samples, d_rgb8, (cudaStream_t) stream);
    RTB_CUDA(cudaGetLastError());
    return RTB_OK;
}

}  // extern "C"
