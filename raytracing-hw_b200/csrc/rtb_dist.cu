// Multi-GPU render: the partition of Scene::render (src/core/scene.cpp:17-65) over the GPUs of
// one box and the ONE exchange step a frame has -- the combine of the per-GPU results on rank 0 over NCCL / NVLink.
// The reference has no distributed mode; the seam is main()'s `render()` then `draw_into()` (src/main.cpp:53-57).
//
//   RTB_SPLIT_TILE   : rows in interleaved blocks, block b -> rank b % N.  Every pixel keeps its own sequential engine
//                      stream, so the image is bit-identical to the one-GPU image.  Combine: every rank ncclSend()s its
//                      finished RGB8 rows (compact) to rank 0, which ncclRecv()s them side by side and puts the rows
//                      in place with one kernel.
//   RTB_SPLIT_SAMPLE : every rank renders the whole canvas with its share of K >= N engine streams per pixel
//                      (rtb_render_params::substreams).  Combine: ncclReduce(sum, fp32, W*H*3) of the partial sums to
//                      rank 0 (NVLS in-switch reduction when NCCL offers it), then the resolve (x 1/samples, ACES,
//                      gamma, RGB8: src/core/scene.cpp:54-63) on rank 0 behind it on the same stream.
//
// One rank per GPU: a process per GPU (bench.py under torchrun) or a thread per GPU (the CLI, rtb_render_multi) -- the
// same code.  NCCL is loaded at run time (dlopen "libnccl.so.2": the copy a hosting torch process has already loaded,
// else the system's), so librtb.so itself links nothing but the CUDA runtime.
#include "../../include/rtb.h"
#include "rtb_kernels.cuh"
#include "rtb_multisync.h"

#include <dlfcn.h>
#include <algorithm>
#include <cstring>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

namespace rtb {
void set_error(const std::string &e);
int scene_scratch(rtb_scene *sc, int which, size_t bytes, void **out);
void blocks_defer_frees(bool on);
void blocks_flush_deferred();
cudaStream_t scene_stream(rtb_scene *sc);
}  // namespace rtb

namespace {

// the few NCCL entry points used, by their published C signatures (nccl.h, stable across 2.x)
struct NcclId { char internal[RTB_UNIQUE_ID_BYTES]; };
typedef struct ncclComm *ncclComm_p;
enum { kNcclUint8 = 1, kNcclFloat32 = 7, kNcclSum = 0 };
struct NcclApi {
    void *handle = nullptr;
    int (*GetVersion)(int *) = nullptr;
    int (*GetUniqueId)(NcclId *) = nullptr;
    int (*CommInitRank)(ncclComm_p *, int, NcclId, int) = nullptr;
    int (*CommDestroy)(ncclComm_p) = nullptr;
    int (*Reduce)(const void *, void *, size_t, int, int, int, ncclComm_p, cudaStream_t) = nullptr;
    int (*Send)(const void *, size_t, int, int, ncclComm_p, cudaStream_t) = nullptr;
    int (*Recv)(void *, size_t, int, int, ncclComm_p, cudaStream_t) = nullptr;
    int (*GroupStart)() = nullptr;
    int (*GroupEnd)() = nullptr;
    const char *(*GetErrorString)(int) = nullptr;
    std::string error;
};

NcclApi *nccl() {
    static NcclApi api;
    static std::once_flag once;
    std::call_once(once, [] {
        const char *names[] = {std::getenv("RTB_NCCL_LIB"), "libnccl.so.2", "libnccl.so"};
        for (const char *n : names) {
            if (!n || !*n) continue;
            api.handle = dlopen(n, RTLD_NOW | RTLD_LOCAL);
            if (api.handle) break;
        }
        if (!api.handle) { api.error = std::string("NCCL not found (dlopen libnccl.so.2): ") + (dlerror() ? dlerror() : ""); return; }
        bool ok = true;
        auto sym = [&](const char *name) { void *p = dlsym(api.handle, name); if (!p) { ok = false; api.error = std::string("NCCL symbol missing: ") + name; } return p; };
        api.GetVersion = (int (*)(int *)) sym("ncclGetVersion");
        api.GetUniqueId = (int (*)(NcclId *)) sym("ncclGetUniqueId");
        api.CommInitRank = (int (*)(ncclComm_p *, int, NcclId, int)) sym("ncclCommInitRank");
        api.CommDestroy = (int (*)(ncclComm_p)) sym("ncclCommDestroy");
        api.Reduce = (int (*)(const void *, void *, size_t, int, int, int, ncclComm_p, cudaStream_t)) sym("ncclReduce");
        api.Send = (int (*)(const void *, size_t, int, int, ncclComm_p, cudaStream_t)) sym("ncclSend");
        api.Recv = (int (*)(void *, size_t, int, int, ncclComm_p, cudaStream_t)) sym("ncclRecv");
        api.GroupStart = (int (*)()) sym("ncclGroupStart");
        api.GroupEnd = (int (*)()) sym("ncclGroupEnd");
        api.GetErrorString = (const char *(*)(int)) sym("ncclGetErrorString");
        if (!ok) { dlclose(api.handle); api.handle = nullptr; }
    });
    return &api;
}

thread_local bool t_nccl_failed = false;   // an NCCL call failed on this thread (rtb_render_multi: the communicators are not kept)
int nccl_fail(const char *what, int rc) {
    NcclApi *a = nccl();
    t_nccl_failed = true;
    rtb::set_error(std::string(what) + ": " + (a->GetErrorString ? a->GetErrorString(rc) : "NCCL error"));
    return RTB_ERR_CUDA;
}
#define RTB_NCCL(call)                                   \
    do {                                                 \
        const int rc_ = (call);                          \
        if (rc_ != 0) return nccl_fail(#call, rc_);      \
    } while (0)
#define RTB_CU(call)                                                                           \
    do {                                                                                       \
        const cudaError_t e_ = (call);                                                         \
        if (e_ != cudaSuccess) { rtb::set_error(std::string(#call) + ": " + cudaGetErrorString(e_)); return RTB_ERR_CUDA; } \
    } while (0)

// rows of the ranks' compact blocks -> their places in the frame.  src = the ranks' blocks side by side (rank r at
// element offset off[r] rows); one thread per 4 bytes of a row where the row length allows, else per byte.
struct RowMap {
    int width_bytes, height, count, rows_per_block;
    unsigned long long row_off[64];   // first row of rank r's block in src
};
__global__ void __launch_bounds__(256) k_place_rows(const uint8_t *__restrict__ src, uint8_t *__restrict__ dst, const RowMap m) {
    const size_t i = (size_t) blockIdx.x * blockDim.x + threadIdx.x;   // byte of the frame
    const size_t total = (size_t) m.width_bytes * (size_t) m.height;
    if (i >= total) return;
    const int row = (int) (i / (size_t) m.width_bytes), col = (int) (i % (size_t) m.width_bytes);
    const int blk = row / m.rows_per_block, r = blk % m.count;
    // local row of `row` on rank r: full blocks of that rank before it + the row inside its block
    const int local = (blk / m.count) * m.rows_per_block + row % m.rows_per_block;
    dst[i] = src[(m.row_off[r] + (unsigned long long) local) * (unsigned long long) m.width_bytes + col];
}

}  // namespace

struct rtb_comm {
    ncclComm_p comm = nullptr;
    int world = 1, rank = 0, device = 0;
    MultiSync *sync = nullptr;   // set by rtb_render_multi for the duration of a frame
};

extern "C" {

int rtb_comm_unique_id(void *id) {
    if (!id) { rtb::set_error("rtb_comm_unique_id: null argument"); return RTB_ERR_INVALID; }
    NcclApi *a = nccl();
    if (!a->handle) { rtb::set_error(a->error); return RTB_ERR_NO_DEVICE; }
    RTB_NCCL(a->GetUniqueId((NcclId *) id));
    return RTB_OK;
}

int rtb_comm_create(const void *id, int world, int rank, int device, rtb_comm **out) {
    if (!out || world < 1 || world > 64 || rank < 0 || rank >= world || (world > 1 && !id)) { rtb::set_error("rtb_comm_create: bad argument"); return RTB_ERR_INVALID; }
    *out = nullptr;
    if (rtb_device_count() <= 0) { rtb::set_error("no CUDA device available (this library has no CPU fallback)"); return RTB_ERR_NO_DEVICE; }
    rtb_comm *c = new rtb_comm();
    c->world = world; c->rank = rank; c->device = device;
    if (world > 1) {
        NcclApi *a = nccl();
        if (!a->handle) { rtb::set_error(a->error); delete c; return RTB_ERR_NO_DEVICE; }
        if (cudaSetDevice(device) != cudaSuccess) { cudaGetLastError(); rtb::set_error("rtb_comm_create: bad device"); delete c; return RTB_ERR_INVALID; }
        NcclId nid;
        std::memcpy(&nid, id, sizeof nid);
        const int rc = a->CommInitRank(&c->comm, world, nid, rank);
        if (rc != 0) { delete c; return nccl_fail("ncclCommInitRank", rc); }
    }
    *out = c;
    return RTB_OK;
}

void rtb_comm_destroy(rtb_comm *c) {
    if (!c) return;
    if (c->comm) { cudaSetDevice(c->device); nccl()->CommDestroy(c->comm); }
    delete c;
}
int rtb_comm_world(const rtb_comm *c) { return c ? c->world : 1; }
int rtb_comm_rank(const rtb_comm *c) { return c ? c->rank : 0; }

int rtb_dist_rank_params(const rtb_render_params *frame, int split, int world, int rank, rtb_render_params *out) {
    if (!frame || !out || world < 1 || rank < 0 || rank >= world || (split != RTB_SPLIT_TILE && split != RTB_SPLIT_SAMPLE)) {
        rtb::set_error("rtb_dist_rank_params: bad argument");
        return RTB_ERR_INVALID;
    }
    *out = *frame;
    out->row_begin = 0;
    out->row_end = frame->height;
    out->sample_begin = 0;
    if (split == RTB_SPLIT_TILE) {
        out->interleave_count = world > 1 ? world : 0;
        out->interleave_index = world > 1 ? rank : 0;
        out->interleave_rows = frame->interleave_rows > 0 ? frame->interleave_rows : 8;
        if (world <= 1) out->interleave_rows = 0;
    } else {
        out->interleave_count = out->interleave_index = out->interleave_rows = 0;
        const int K = std::max(std::max(1, frame->substreams), world);
        out->substreams = K;
        out->substream_begin = (int) ((long long) K * rank / world);
        out->substream_end = (int) ((long long) K * (rank + 1) / world);
    }
    return RTB_OK;
}

int rtb_render_dist_device(rtb_scene *scene, rtb_comm *comm, const rtb_render_params *frame, int split, float *d_accum_root,
                           uint8_t *d_rgb8_root, unsigned long long *d_counters, void *stream_) {
    if (!scene || !frame) { rtb::set_error("rtb_render_dist_device: null argument"); return RTB_ERR_INVALID; }
    const int world = comm ? comm->world : 1, rank = comm ? comm->rank : 0;
    cudaStream_t stream = (cudaStream_t) stream_;
    rtb_render_params p;
    int rc = rtb_dist_rank_params(frame, split, world, rank, &p);
    if (rc != RTB_OK) return rc;
    if (comm && comm->device != rtb_scene_device(scene)) { rtb::set_error("rtb_render_dist_device: the scene and the communicator live on different devices"); return RTB_ERR_INVALID; }
    RTB_CU(cudaSetDevice(rtb_scene_device(scene)));
    const size_t W = (size_t) frame->width, H = (size_t) frame->height;
    NcclApi *a = world > 1 ? nccl() : nullptr;

    if (split == RTB_SPLIT_SAMPLE) {
        const bool whole = p.substream_end - p.substream_begin == p.substreams;   // one rank holds every stream
        if (whole) return rtb_render_device(scene, &p, d_accum_root, d_rgb8_root, d_counters, stream_);
        void *local = nullptr, *total = nullptr;
        rc = rtb::scene_scratch(scene, 0, W * H * 3 * sizeof(float), &local);
        if (rc == RTB_OK) rc = rtb_render_device(scene, &p, (float *) local, nullptr, d_counters, stream_);
        if (rc == RTB_OK && rank == 0) {
            total = d_accum_root;
            if (!total) rc = rtb::scene_scratch(scene, 1, W * H * 3 * sizeof(float), &total);
        }
        if (comm && comm->sync) {   // all ranks of this process enter the collective, or none
            const int all = comm->sync->arrive(rank, rc);
            if (rc == RTB_OK && all != RTB_OK) rtb::set_error("rtb_render_dist_device: another rank of the frame failed");
            if (all != RTB_OK) return rc != RTB_OK ? rc : all;
        } else if (rc != RTB_OK) {
            return rc;
        }
        RTB_NCCL(a->Reduce(local, total, W * H * 3, kNcclFloat32, kNcclSum, 0, comm->comm, stream));
        if (rank == 0 && d_rgb8_root) rtb::launch_resolve((const float *) total, W * H, frame->samples, d_rgb8_root, stream);
        RTB_CU(cudaGetLastError());
        return RTB_OK;
    }

    // ---- tile split
    if (world == 1) return rtb_render_device(scene, &p, d_accum_root, d_rgb8_root, d_counters, stream_);
    if (d_accum_root) { rtb::set_error("rtb_render_dist_device: the tile split gathers the finished RGB8 rows only (accum must be NULL)"); return RTB_ERR_INVALID; }
    const int local_rows = rtb_render_local_rows(&p);
    if (local_rows < 0) { rtb::set_error("rtb_render_dist_device: bad canvas"); return RTB_ERR_INVALID; }
    RowMap m;
    m.width_bytes = (int) (W * 3); m.height = (int) H; m.count = world; m.rows_per_block = p.interleave_rows;
    std::vector<int> rows((size_t) world);
    unsigned long long off = 0;
    for (int r = 0; r < world; ++r) {
        rtb_render_params q = p;
        q.interleave_index = r;
        rows[(size_t) r] = rtb_render_local_rows(&q);
        m.row_off[r] = off;
        off += (unsigned long long) rows[(size_t) r];
    }
    void *local = nullptr, *side = nullptr;
    if (rank == 0) {
        // rank 0 renders straight into its place in the side-by-side buffer (offset 0)
        rc = rtb::scene_scratch(scene, 1, (size_t) off * W * 3 + 256, &side);
        local = side;
    } else {
        rc = rtb::scene_scratch(scene, 0, (size_t) local_rows * W * 3 + 256, &local);
    }
    if (rc == RTB_OK) rc = rtb_render_device(scene, &p, nullptr, (uint8_t *) local, d_counters, stream_);
    if (comm && comm->sync) {   // all ranks of this process enter the collective, or none
        const int all = comm->sync->arrive(rank, rc);
        if (rc == RTB_OK && all != RTB_OK) rtb::set_error("rtb_render_dist_device: another rank of the frame failed");
        if (all != RTB_OK) return rc != RTB_OK ? rc : all;
    } else if (rc != RTB_OK) {
        return rc;
    }
    RTB_NCCL(a->GroupStart());
    int xfer = 0;   // first failure inside the group; the group is closed whatever happens (an open group swallows the thread's next NCCL calls)
    if (rank == 0) {
        for (int r = 1; r < world && xfer == 0; ++r)
            if (rows[(size_t) r] > 0) xfer = a->Recv((uint8_t *) side + m.row_off[r] * W * 3, (size_t) rows[(size_t) r] * W * 3, kNcclUint8, r, comm->comm, stream);
    } else if (local_rows > 0) {
        xfer = a->Send(local, (size_t) local_rows * W * 3, kNcclUint8, 0, comm->comm, stream);
    }
    const int ended = a->GroupEnd();
    if (xfer != 0) return nccl_fail(rank == 0 ? "ncclRecv" : "ncclSend", xfer);
    if (ended != 0) return nccl_fail("ncclGroupEnd", ended);
    if (rank == 0 && d_rgb8_root) {
        const size_t total = W * 3 * H;
        k_place_rows<<<(unsigned) ((total + 255) / 256), 256, 0, stream>>>((const uint8_t *) side, d_rgb8_root, m);
    }
    RTB_CU(cudaGetLastError());
    return RTB_OK;
}

int rtb_render_dist(rtb_scene *scene, rtb_comm *comm, const rtb_render_params *frame, int split, uint8_t *rgb8_root, rtb_stats *stats) {
    if (!scene || !frame) { rtb::set_error("rtb_render_dist: null argument"); return RTB_ERR_INVALID; }
    if (stats) std::memset(stats, 0, sizeof(*stats));
    RTB_CU(cudaSetDevice(rtb_scene_device(scene)));
    const int rank = comm ? comm->rank : 0;
    const size_t bytes = (size_t) frame->width * (size_t) frame->height * 3;
    void *buf = nullptr;
    int rc = rtb::scene_scratch(scene, 2, bytes + 256 + 64, &buf);
    if (rc != RTB_OK) return rc;
    uint8_t *d_rgb = (uint8_t *) buf;
    unsigned long long *d_cnt = (unsigned long long *) ((char *) buf + (bytes + 255) / 256 * 256);
    cudaStream_t st = rtb::scene_stream(scene);
    RTB_CU(cudaMemsetAsync(d_cnt, 0, 48, st));
    cudaEvent_t e0, e1;
    RTB_CU(cudaEventCreate(&e0));
    RTB_CU(cudaEventCreate(&e1));
    const unsigned long long launches0 = rtb_scene_launches(scene);
    cudaEventRecord(e0, st);
    rc = rtb_render_dist_device(scene, comm, frame, split, nullptr, rank == 0 ? d_rgb : nullptr, d_cnt, st);
    cudaEventRecord(e1, st);
    unsigned long long c[6] = {0, 0, 0, 0, 0, 0};
    if (rc == RTB_OK) {
        if (rank == 0 && rgb8_root) cudaMemcpyAsync(rgb8_root, d_rgb, bytes, cudaMemcpyDeviceToHost, st);
        cudaMemcpyAsync(c, d_cnt, 48, cudaMemcpyDeviceToHost, st);
        const cudaError_t e = cudaStreamSynchronize(st);
        if (e != cudaSuccess) { rtb::set_error(std::string("rtb_render_dist: ") + cudaGetErrorString(e)); rc = RTB_ERR_CUDA; }
    }
    if (rc == RTB_OK && stats) {
        float ms = 0.f;
        cudaEventElapsedTime(&ms, e0, e1);
        stats->kernel_ms = ms;
        stats->closest_hit_rays = c[0];
        stats->light_pdf_queries = c[1];
        stats->box_tests = c[2]; stats->tri_tests = c[3];
        stats->light_box_tests = c[4]; stats->light_tri_tests = c[5];
        stats->launches = rtb_scene_launches(scene) - launches0;
    }
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    return rc;
}

// One process, one thread per device: what the CLI runs.  scenes[r] lives on the device rank r renders on.  The
// communicators of a device list are created once (ncclCommInitRank of 8 ranks costs about a frame) and kept.
int rtb_render_multi(rtb_scene *const *scenes, int n, const rtb_render_params *frame, int split, uint8_t *rgb8, rtb_stats *stats) {
    if (!scenes || n < 1 || n > 64 || !frame) { rtb::set_error("rtb_render_multi: bad argument"); return RTB_ERR_INVALID; }
    if (stats) std::memset(stats, 0, sizeof(*stats));
    if (n == 1) return rtb_render_dist(scenes[0], nullptr, frame, split, rgb8, stats);
    static std::mutex mu;
    static std::vector<std::pair<std::vector<int>, std::vector<rtb_comm *>>> kept;
    std::lock_guard<std::mutex> lk(mu);   // (one multi-device frame at a time per process)
    std::vector<int> devs((size_t) n);
    for (int r = 0; r < n; ++r) devs[(size_t) r] = rtb_scene_device(scenes[r]);
    std::vector<rtb_comm *> *comms = nullptr;
    for (int r = 0; r < n; ++r)
        for (int q = 0; q < r; ++q)
            if (devs[(size_t) r] == devs[(size_t) q]) { rtb::set_error("rtb_render_multi: two scenes on device " + std::to_string(devs[(size_t) r]) + " (one rank per device)"); return RTB_ERR_INVALID; }
    for (auto &k : kept) if (k.first == devs) comms = &k.second;
    std::vector<int> rcs((size_t) n, RTB_OK);
    std::vector<std::string> errs((size_t) n);
    std::vector<rtb_stats> st((size_t) n);
    char id[RTB_UNIQUE_ID_BYTES];
    std::vector<rtb_comm *> fresh;
    if (!comms) {
        const int rc = rtb_comm_unique_id(id);
        if (rc != RTB_OK) return rc;
        fresh.assign((size_t) n, nullptr);
    }
    MultiSync sync, presync;   // before the collective of the frame / before ncclCommInitRank (which waits for all n ranks)
    sync.n = n;
    presync.n = n;
    std::vector<char> nccl_bad((size_t) n, 0);
    std::vector<std::thread> th;
    for (int r = 0; r < n; ++r) {
        th.emplace_back([&, r] {
            rtb::blocks_defer_frees(true);   // no cudaFree on a thread that holds a rank (BlockCache::release)
            int e = RTB_OK;
            rtb_comm *c = comms ? (*comms)[(size_t) r] : nullptr;
            t_nccl_failed = false;
            if (!comms) {
                // what can fail on ONE rank only is tried first and the ranks compare notes: a rank that cannot reach its device
                // must not leave the other n - 1 waiting in ncclCommInitRank for ever
                int pre = RTB_OK;
                if (cudaSetDevice(devs[(size_t) r]) != cudaSuccess) {
                    cudaGetLastError();
                    rtb::set_error("rtb_render_multi: device " + std::to_string(devs[(size_t) r]) + " cannot be selected");
                    pre = RTB_ERR_CUDA;
                }
                const int all = presync.arrive(r, pre);
                if (pre == RTB_OK && all != RTB_OK) rtb::set_error("rtb_render_multi: another rank could not initialise");
                e = pre != RTB_OK ? pre : all;
                if (e == RTB_OK) e = rtb_comm_create(id, n, r, devs[(size_t) r], &c);
                fresh[(size_t) r] = c;
            }
            // (a rank without a communicator cannot take part: it still has to be counted at the meeting point)
            if (c) c->sync = &sync;
            if (e == RTB_OK) e = rtb_render_dist(scenes[r], c, frame, split, r == 0 ? rgb8 : nullptr, &st[(size_t) r]);
            if (!sync.met[r]) sync.arrive(r, e);   // returned before the meeting point: the others must not wait
            if (c) c->sync = nullptr;
            if (e != RTB_OK) errs[(size_t) r] = rtb_last_error();
            nccl_bad[(size_t) r] = t_nccl_failed ? 1 : 0;
            rcs[(size_t) r] = e;
        });
    }
    for (auto &t : th) t.join();
    rtb::blocks_flush_deferred();
    bool nccl_failed = false;
    for (char b : nccl_bad) nccl_failed = nccl_failed || b != 0;
    if (!comms) {
        bool all = !nccl_failed;
        for (auto *c : fresh) all = all && c != nullptr;
        if (all) kept.emplace_back(devs, fresh);
        else for (auto *c : fresh) rtb_comm_destroy(c);
    } else if (nccl_failed) {
        // a communicator that has reported an error is not used again: the next frame creates a new set
        for (size_t k = 0; k < kept.size(); ++k)
            if (&kept[k].second == comms) {
                for (auto *c : kept[k].second) rtb_comm_destroy(c);
                kept.erase(kept.begin() + (long) k);
                break;
            }
        comms = nullptr;
    }
    for (int r = 0; r < n; ++r)
        if (rcs[(size_t) r] != RTB_OK) {
            // a rank thread that ran out of memory could not trim the block cache itself (BlockCache::acquire): do it here, where
            // no thread holds a rank, so that calling again has the cached blocks' memory to work with
            rtb_release_cached_memory();
            rtb::set_error("rank " + std::to_string(r) + ": " + errs[(size_t) r]);
            return rcs[(size_t) r];
        }
    if (stats)
        for (int r = 0; r < n; ++r) {
            stats->closest_hit_rays += st[(size_t) r].closest_hit_rays;
            stats->light_pdf_queries += st[(size_t) r].light_pdf_queries;
            stats->box_tests += st[(size_t) r].box_tests; stats->tri_tests += st[(size_t) r].tri_tests;
            stats->light_box_tests += st[(size_t) r].light_box_tests; stats->light_tri_tests += st[(size_t) r].light_tri_tests;
            stats->launches += st[(size_t) r].launches;
            stats->kernel_ms = std::max(stats->kernel_ms, st[(size_t) r].kernel_ms);
        }
    return RTB_OK;
}

}  // extern "C"
