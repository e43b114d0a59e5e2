// Wavefront integrator (sm_100a): Scene::render (src/core/scene.cpp:17-65) as a ray queue.
//
//   k_wf_init    : one thread per slot (pixel): engine seed, first jitter, first primary ray     (scene.cpp:34-39)
//   k_wf_trace   : persistent warps; lanes pull the active slots' rays from a queue and run the voted node /
//                  triangle phases of rtb_trace.cuh (BVH::intersect, src/core/bvh.cpp:177-243)
//   k_wf_shade   : one thread per active slot: Scene::intersect's vertex (scene.cpp:85-154) on the hit just found,
//                  next direction from the slot's own sample stream, fold + accumulate + next sample at a path's end
//   k_wf_count / k_wf_scan / k_wf_scatter : ORDER-PRESERVING compaction of the active list (finished pixels leave;
//                  neighbouring pixels stay neighbours, so state loads stay coalesced)
//
// The host enqueues samples x ray_depth iterations (the most closest-hit rays any pixel can need) without ever
// synchronising: every kernel reads the live count of its list from device memory, so iterations after the last
// pixel finished cost an empty launch.  See rtb_wavefront.cuh for the state layout and the ncu evidence.
#include "rtb_wavefront.cuh"

#include <algorithm>
#include <chrono>
#include <cstdlib>
#include <thread>
#include <utility>
#include <vector>
#include "rtb_trace.cuh"

namespace rtb {

namespace {
constexpr int kShadeBlock = 128;
constexpr int kDefaultShadeBlocks = 6;   // CTAs per SM the shade kernel is compiled for (RTB_SHADE_BLOCKS: 5, 6, 8)
constexpr int kCompactBlock = 1024;
constexpr int kDefaultTailSteps = 96;      // see trace_persistent (RTB_TRACE_TAIL; 0 = rays are never suspended)
constexpr int kDefaultRaysPerLane = 1;   // k_wf_trace keeps ceil(active / (128 * this)) CTAs (RTB_TRACE_RPL)

// slots [slot_begin, slot_begin + count) of one slot group; list / n_list / queue_head are the group's own
__global__ void __launch_bounds__(256) k_wf_init(const __grid_constant__ DScene sc, const RenderLaunch p, const WfState w, unsigned slot_begin,
                                                 unsigned count, unsigned *__restrict__ list, unsigned *__restrict__ n_list,
                                                 unsigned *__restrict__ queue_head) {
    const unsigned i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i == 0) {
        *n_list = count;
        *queue_head = 0u;
    }
    if (i < count) list[i] = wf_init_slot(sc, p, w, slot_begin + i);
}

template <bool COUNT, int MINB, int MODE>
__global__ void __launch_bounds__(128, MINB) k_wf_trace(const __grid_constant__ DScene sc, const WfState w, unsigned *list,
                                                  const unsigned *__restrict__ n_list, unsigned *__restrict__ queue_head,
                                                  unsigned long long *__restrict__ counters, const TracePolicy pol,
                                                  const unsigned rays_per_lane, const int tail_steps) {
    const unsigned n = *n_list;
    // the grid was sized on the host for the list's capacity; when few slots are still active, the surplus CTAs leave
    // at once so that every remaining lane still pulls several rays from the queue (dynamic balance needs that)
    if (n == 0 || (blockIdx.x > 0 && blockIdx.x * (128u * rays_per_lane) >= n)) return;
    QueueCounters cnt = {0ull, 0ull, 0u};
    // the tag of a ray is its index in the active list: the entry says which slot it is and whether it was suspended
    trace_persistent<COUNT, MODE, RTB_RING_WAVEFRONT>(
        sc.bvh, n, queue_head,
        [&](unsigned i, LaneTraversal<COUNT, RTB_RING_WAVEFRONT> &L, unsigned &tag) {
            const unsigned e = list[i];
            if (e & kSlotDone) return 0;
            tag = i;
            const unsigned slot = e & kSlotIndex;
#if RTB_TRACE_STATE_CS
            {
                const float4 ra = __ldcs(w.ray_a + slot);
                const float2 rb = __ldcs(w.ray_b + slot);
                make_ray(L.r, ra.x, ra.y, ra.z, ra.w, rb.x, rb.y);
            }
#else
            wf_load_ray(w, slot, L.r);
#endif
            if (e & kSlotSuspended) {
                L.load3(w.susp + (size_t) slot * w.susp_stride);
                return 2;
            }
            return 1;
        },
        [&](unsigned i, const Ray &, const Hit &h) {
            const unsigned e = list[i];
#if RTB_TRACE_STATE_CS
            __stcs(w.hit + (e & kSlotIndex), make_float4(h.t, h.u, h.v, u2f((uint32_t) h.id)));
#else
            w.hit[e & kSlotIndex] = make_float4(h.t, h.u, h.v, u2f((uint32_t) h.id));
#endif
            if (e & kSlotSuspended) list[i] = e & ~kSlotSuspended;
        },
        [&](unsigned i, LaneTraversal<COUNT, RTB_RING_WAVEFRONT> &L) {
            const unsigned e = list[i];
            L.save3(w.susp + (size_t) (e & kSlotIndex) * w.susp_stride);
            if (!(e & kSlotSuspended)) list[i] = e | kSlotSuspended;
        },
        tail_steps, &cnt, pol);
    if (counters) {
        for (int o = 16; o > 0; o >>= 1) {
            cnt.rays += __shfl_down_sync(0xffffffffu, cnt.rays, o);
            if (COUNT) {
                cnt.box += __shfl_down_sync(0xffffffffu, cnt.box, o);
                cnt.tri += __shfl_down_sync(0xffffffffu, cnt.tri, o);
            }
        }
        if ((threadIdx.x & 31) == 0) {
            if (cnt.rays) atomicAdd(&counters[0], (unsigned long long) cnt.rays);
            if (COUNT) {
                atomicAdd(&counters[2], cnt.box);
                atomicAdd(&counters[3], cnt.tri);
            }
        }
    }
}

template <bool COUNT, int MINB>
__global__ void __launch_bounds__(kShadeBlock, MINB) k_wf_shade(const __grid_constant__ DScene sc, const RenderLaunch p, const WfState w,
                                                          unsigned *__restrict__ list, const unsigned *__restrict__ n_list,
                                                          unsigned *__restrict__ queue_head) {
    const unsigned n = *n_list;
    if (blockIdx.x == 0 && threadIdx.x == 0) *queue_head = 0u;   // the next k_wf_trace starts at the head of the list
    PathCounters pc = {0ull, 0ull, {0ull, 0ull}, {0ull, 0ull}};
    for (unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const unsigned e = list[i];
        if (e & (kSlotDone | kSlotSuspended)) continue;   // (a suspended ray has no hit yet: the slot waits for a later launch)
        const unsigned ne = wf_shade_slot<COUNT>(sc, p, w, e, &pc);
        if (ne != e) list[i] = ne;
    }
    if (p.d_counters) {
        // d_counters: {rays, light queries, box, tri, light box, light tri}; box / tri are totals over both BVHs
        unsigned long long c[3] = {pc.light_queries, pc.light.box, pc.light.tri};
#pragma unroll
        for (int k = 0; k < (COUNT ? 3 : 1); ++k) {
            for (int o = 16; o > 0; o >>= 1) c[k] += __shfl_down_sync(0xffffffffu, c[k], o);
            if ((threadIdx.x & 31) == 0 && c[k]) {
                atomicAdd(&p.d_counters[k + 1], c[k]);
                if (k > 0) atomicAdd(&p.d_counters[k + 3], c[k]);
            }
        }
    }
}

// sample-range split: one thread per local pixel folds the pixel's streams in order (wf_fold_pixel)
__global__ void __launch_bounds__(256) k_wf_fold(const RenderLaunch p, const WfState w) {
    const size_t i = (size_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (size_t) p.local_rows * (size_t) p.width) return;
    wf_fold_pixel(p, w, (int) (i % (size_t) p.width), (int) (i / (size_t) p.width));
}

// ---- order-preserving compaction: count per block, scan the block counts, scatter
__global__ void __launch_bounds__(kCompactBlock) k_wf_count(const unsigned *__restrict__ list, const unsigned *__restrict__ n_list,
                                                            unsigned *__restrict__ block_counts) {
    __shared__ unsigned warp_sums[kCompactBlock / 32];
    const unsigned n = *n_list;
    const unsigned i = blockIdx.x * kCompactBlock + threadIdx.x;
    const bool alive = i < n && !(list[i] & kSlotDone);
    const unsigned m = __ballot_sync(0xffffffffu, alive);
    if ((threadIdx.x & 31) == 0) warp_sums[threadIdx.x >> 5] = __popc(m);
    __syncthreads();
    if (threadIdx.x < 32) {
        unsigned v = warp_sums[threadIdx.x];
        for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
        if (threadIdx.x == 0) block_counts[blockIdx.x] = v;
    }
}

// one block: exclusive scan of the block counts in place; the total becomes the length of the output list
__global__ void __launch_bounds__(1024) k_wf_scan(unsigned *__restrict__ block_counts, unsigned n_blocks, unsigned *__restrict__ n_out) {
    __shared__ unsigned warp_sums[32];
    __shared__ unsigned carry;
    if (threadIdx.x == 0) carry = 0u;
    __syncthreads();
    for (unsigned base = 0; base < n_blocks; base += 1024) {
        const unsigned i = base + threadIdx.x;
        const unsigned v = i < n_blocks ? block_counts[i] : 0u;
        unsigned inc = v;
        for (int o = 1; o < 32; o <<= 1) {
            const unsigned t = __shfl_up_sync(0xffffffffu, inc, o);
            if ((threadIdx.x & 31) >= (unsigned) o) inc += t;
        }
        if ((threadIdx.x & 31) == 31) warp_sums[threadIdx.x >> 5] = inc;
        __syncthreads();
        if (threadIdx.x < 32) {
            unsigned s = warp_sums[threadIdx.x];
            for (int o = 1; o < 32; o <<= 1) {
                const unsigned t = __shfl_up_sync(0xffffffffu, s, o);
                if (threadIdx.x >= (unsigned) o) s += t;
            }
            warp_sums[threadIdx.x] = s;   // inclusive over warps
        }
        __syncthreads();
        const unsigned warp_off = (threadIdx.x >> 5) ? warp_sums[(threadIdx.x >> 5) - 1] : 0u;
        const unsigned c = carry;
        if (i < n_blocks) block_counts[i] = c + warp_off + inc - v;
        __syncthreads();
        if (threadIdx.x == 1023) carry = c + warp_off + inc;
        __syncthreads();
    }
    if (threadIdx.x == 0) *n_out = carry;
}

__global__ void __launch_bounds__(kCompactBlock) k_wf_scatter(const unsigned *__restrict__ list, const unsigned *__restrict__ n_list,
                                                              const unsigned *__restrict__ block_offsets, unsigned *__restrict__ out) {
    __shared__ unsigned warp_sums[kCompactBlock / 32];
    const unsigned n = *n_list;
    const unsigned i = blockIdx.x * kCompactBlock + threadIdx.x;
    const unsigned e = i < n ? list[i] : kSlotDone;
    const bool alive = !(e & kSlotDone);
    const unsigned m = __ballot_sync(0xffffffffu, alive);
    const unsigned lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (lane == 0) warp_sums[warp] = __popc(m);
    __syncthreads();
    if (threadIdx.x < 32) {
        unsigned s = warp_sums[threadIdx.x];
        for (int o = 1; o < 32; o <<= 1) {
            const unsigned t = __shfl_up_sync(0xffffffffu, s, o);
            if (threadIdx.x >= (unsigned) o) s += t;
        }
        warp_sums[threadIdx.x] = s;
    }
    __syncthreads();
    if (alive) {
        const unsigned off = block_offsets[blockIdx.x] + (warp ? warp_sums[warp - 1] : 0u) + __popc(m & ((1u << lane) - 1u));
        out[off] = e;
    }
}
}  // namespace

static unsigned wf_susp_stride(int bvh_depth) { return (unsigned) (LaneTraversal<false>::kSuspHeader + std::max(0, bvh_depth) + 3); }

size_t wf_state_bytes(unsigned n_slots, int depth_cap, int bvh_depth) {
    const size_t per_slot = 16 + 8 + 16 + 16 + 16 + (size_t) depth_cap * (16 + 16 + 4) + 2 * 4 + (size_t) wf_susp_stride(bvh_depth) * 8;
    const size_t blocks = ((size_t) n_slots + kCompactBlock - 1) / kCompactBlock + kWfMaxGroups;
    return per_slot * n_slots + blocks * 4 + 4096;
}

// Carves the state arrays out of one allocation (`base`, wf_alloc_bytes() bytes, 256-byte aligned).
static void wf_carve(char *base, unsigned n_slots, int depth_cap, int bvh_depth, WfState &w, unsigned *lists[2], unsigned **block_counts, unsigned **scalars) {
    size_t off = 0;
    auto take = [&](size_t bytes) {
        char *p = base + off;
        off += (bytes + 255) / 256 * 256;
        return p;
    };
    w.n_slots = n_slots;
    w.depth_cap = depth_cap;
    w.ray_a = (float4 *) take((size_t) n_slots * 16);
    w.hit = (float4 *) take((size_t) n_slots * 16);
    w.ctl = (uint4 *) take((size_t) n_slots * 16);
    w.sum = (float4 *) take((size_t) n_slots * 16);
    w.stk_e = (float4 *) take((size_t) n_slots * 16 * depth_cap);
    w.stk_m = (float4 *) take((size_t) n_slots * 16 * depth_cap);
    w.ray_b = (float2 *) take((size_t) n_slots * 8);
    w.stk_a = (float *) take((size_t) n_slots * 4 * depth_cap);
    w.susp_stride = wf_susp_stride(bvh_depth);
    w.susp = (int2 *) take((size_t) n_slots * w.susp_stride * 8);
    lists[0] = (unsigned *) take((size_t) n_slots * 4);
    lists[1] = (unsigned *) take((size_t) n_slots * 4);
    *block_counts = (unsigned *) take((((size_t) n_slots + kCompactBlock - 1) / kCompactBlock + kWfMaxGroups) * 4);
    *scalars = (unsigned *) take(256);
}

size_t wf_alloc_bytes(unsigned n_slots, int depth_cap, int bvh_depth) {
    // every array is rounded up to 256 bytes by wf_carve: 13 arrays
    return wf_state_bytes(n_slots, depth_cap, bvh_depth) + 13 * 256;
}

unsigned wf_plane_slots(const RenderLaunch &p) {
    const long long tiles = (long long) ((p.width + kWfTileW - 1) / kWfTileW) * ((p.local_rows + kWfTileH - 1) / kWfTileH);
    return (unsigned) (tiles * 32);
}
unsigned wf_slots(const RenderLaunch &p) { return wf_plane_slots(p) * (unsigned) std::max(1, p.substreams > 1 ? p.sub_end - p.sub_begin : 1); }

// Measured on one B200 (Sponza stand-in, 64 spp, tail suspension on; profiles/r01k_groups_canvas_sweep.txt), Mpix*spp/s:
//   1/2 of a 1080p frame (1.04 M slots): 268 / 258 / 234 with 1 / 2 / 4 groups
//   1/4 (0.52 M slots)                 : 234 / 218 / 216 / 216 with 1 / 2 / 4 / 8 groups
//   1/8 (0.26 M slots)                 : 182 / 194 / 200 / 202 with 2 / 4 / 6 / 8 groups
// (before rays could be suspended, groups paid off from 1/2 frame down: they hid the tails of each other's launches;
// with bounded tails their contention costs more than that until a launch is too small to fill the GPU.)
// Default: one group down to 0.4 M slots, below that as many groups as bring a group down to ~0.26 M / 8 slots.
// Light scenes (a traversal set that sits in the L1s: the 36-triangle boxes of cfg2) are the exception: their queue
// kernel is short and not ALU-bound, the frame is as much shade as trace, and a second group overlaps the two
// (profiles/r01m_groups_light_scene_sweep.txt: 1280x720 590 / 657 / 650 / 636 with 1 / 2 / 3 / 4 groups; 1920x1080
// 605 / 654 / 652 with 1 / 2 / 4; 640x360 461 / 512 / 527 with 1 / 2 / 4).
int wf_groups(unsigned n_slots, bool light_scene) {
    static const int forced = [] {
        int g = 0;
        if (const char *e = std::getenv("RTB_WF_GROUPS")) g = std::max(1, std::min(kWfMaxGroups, std::atoi(e)));
        return g;
    }();
    if (forced) return forced;
    if (n_slots >= 400000u) return light_scene ? 2 : 1;
    const unsigned g = ((2u << 20) + n_slots / 2) / std::max(1u, n_slots);
    return (int) std::max(1u, std::min(light_scene ? 4u : 8u, g));
}

// Slot groups.  The slots are cut into up to kWfMaxGroups contiguous groups, each with its own active list, queue
// cursor and stream (`aux`).  Inside a group the kernels of an iteration are ordered (trace -> shade); the groups are
// independent, so the tail of one group's queue kernel (lanes running dry one by one) and its latency-bound shade
// kernel overlap the other group's ALU-bound queue kernel.  With stage timing on, one group on the caller's stream:
// a kernel's duration is only defined when it runs alone.
int launch_render_wavefront(const DScene &sc, const RenderLaunch &p_in, void *state, cudaStream_t stream, unsigned long long *launches,
                            std::vector<cudaEvent_t> *stage_events, WfStreams *aux) {
    if (p_in.local_rows <= 0 || p_in.width <= 0 || p_in.samples <= 0) return 0;
    int status = 0;
    RenderLaunch p = p_in;
    if (p.substreams <= 1) { p.substreams = 1; p.sub_begin = 0; p.sub_end = 1; }
    p.plane_slots = wf_plane_slots(p);
    const unsigned n_slots = wf_slots(p);
    const int depth_cap = std::max(1, sc.ray_depth - 1);
    WfState w;
    unsigned *lists[2], *block_counts, *scalars;
    wf_carve((char *) state, n_slots, depth_cap, sc.bvh.max_depth, w, lists, &block_counts, &scalars);
    unsigned long long nl = 0;

    // a group needs enough slots to fill the persistent queue kernel a few times over
    const bool light_scene = (size_t) sc.bvh.n_internal * sizeof(DNode) + (size_t) sc.bvh.n_tris * kTriStride * sizeof(float4) <= (64u << 10);
    int G = (stage_events || !aux) ? 1 : wf_groups(n_slots, light_scene);
    while (G > 1 && n_slots / (unsigned) G < 16u * 1024u) --G;
    struct Group {
        unsigned begin, count, n_blocks, block_off;
        unsigned *lists[2], *n_list[2], *queue_head;
        int cur;
        cudaStream_t st;
        unsigned trace_grid, shade_grid;
    } grp[kWfMaxGroups];
    const int bps = persistent_blocks_per_sm();
    {
        const unsigned tiles = n_slots / 32u;
        unsigned t0 = 0, boff = 0;
        for (int g = 0; g < G; ++g) {
            const unsigned t1 = (unsigned) ((unsigned long long) tiles * (g + 1) / G);
            Group &q = grp[g];
            q.begin = t0 * 32u;
            q.count = (t1 - t0) * 32u;
            q.n_blocks = (q.count + kCompactBlock - 1) / kCompactBlock;
            q.block_off = boff;
            boff += q.n_blocks;
            q.lists[0] = lists[0] + q.begin;
            q.lists[1] = lists[1] + q.begin;
            q.n_list[0] = scalars + 4 * g;
            q.n_list[1] = scalars + 4 * g + 1;
            q.queue_head = scalars + 4 * g + 2;
            q.cur = 0;
            q.st = G > 1 ? aux->stream[g] : stream;
            // the groups share the SMs: each gets its share of the persistent grid
            q.trace_grid = (unsigned) std::min<size_t>(((size_t) q.count + 127) / 128, (size_t) persistent_grid(128, bps));
            q.shade_grid = (unsigned) std::min<size_t>(((size_t) q.count + kShadeBlock - 1) / kShadeBlock, (size_t) persistent_grid(kShadeBlock, 16));
            t0 = t1;
        }
    }
    if (G > 1) {   // fork: the groups' streams start after everything already queued on the caller's stream
        cudaEventRecord(aux->fork, stream);
        for (int g = 0; g < G; ++g) cudaStreamWaitEvent(grp[g].st, aux->fork, 0);
    }
    auto compact = [&](Group &q) {
        unsigned *bc = block_counts + q.block_off;
        k_wf_count<<<q.n_blocks, kCompactBlock, 0, q.st>>>(q.lists[q.cur], q.n_list[q.cur], bc);
        k_wf_scan<<<1, 1024, 0, q.st>>>(bc, q.n_blocks, q.n_list[q.cur ^ 1]);
        k_wf_scatter<<<q.n_blocks, kCompactBlock, 0, q.st>>>(q.lists[q.cur], q.n_list[q.cur], bc, q.lists[q.cur ^ 1]);
        q.cur ^= 1;
        nl += 3;
    };
    for (int g = 0; g < G; ++g) {
        Group &q = grp[g];
        k_wf_init<<<(q.count + 255) / 256, 256, 0, q.st>>>(sc, p, w, q.begin, q.count, q.lists[0], q.n_list[0], q.queue_head);
        ++nl;
        compact(q);   // drops the slots outside the canvas
    }
    const TracePolicy pol = trace_policy();
    const int mode = trace_mode();
    static const int shade_bps = [] {
        int v = kDefaultShadeBlocks;
        if (const char *e = std::getenv("RTB_SHADE_BLOCKS")) v = std::atoi(e);
        return v;
    }();
    static const unsigned rays_per_lane = [] {
        int v = kDefaultRaysPerLane;
        if (const char *e = std::getenv("RTB_TRACE_RPL")) v = std::max(1, std::atoi(e));
        return (unsigned) v;
    }();
    // RTB_TRACE_TAIL: steps a ray may run in one launch once the queue is empty before it is suspended (0: never)
    static const int tail_steps_env = [] {
        int v = kDefaultTailSteps;
        if (const char *e = std::getenv("RTB_TRACE_TAIL")) v = std::max(0, std::atoi(e));
        return v;
    }();
    const int tail_steps = p.count_tests ? 0 : tail_steps_env;
    // samples x ray_depth iterations trace every ray a pixel can need when no ray is ever suspended; with suspension a
    // slot sometimes spends an iteration without finishing its ray, so a margin is enqueued up front and the live counts
    // are then read back (the one host synchronisation of a frame) until every group's list is empty
    // (with K streams per pixel the longest stream has ceil(samples / K) samples)
    const long long chain_samples = p.substreams > 1 ? ((long long) p.samples + p.substreams - 1) / p.substreams : (long long) (p.samples - p.sample_begin);
    const long long base_iterations = chain_samples * sc.ray_depth;
    const long long iterations = tail_steps > 0 ? base_iterations + base_iterations / 16 + 8 : base_iterations;
    const int period = 16;   // iterations between compactions
    // optional stage timing: events[2*it], events[2*it+1] bracket iteration it's k_wf_trace; events[2*iterations] ends the frame
    size_t ev_i = 0;
    auto mark = [&]() {
        if (!stage_events) return;
        if (ev_i == stage_events->size()) {
            cudaEvent_t e;
            cudaEventCreate(&e);
            stage_events->push_back(e);
        }
        cudaEventRecord((*stage_events)[ev_i++], stream);
    };
    // Progress line (rtb_scene_set_progress): events on the first group's stream every ~1 % of the planned iterations;
    // once everything is enqueued the host polls them in order and calls back as they complete.
    int (*const progress)(int, void *) = (aux && !stage_events) ? aux->progress : nullptr;
    void *const progress_user = aux ? aux->progress_user : nullptr;
    std::vector<std::pair<cudaEvent_t, int>> prog_ev;
    const long long prog_stride = std::max<long long>(1, iterations / 100);
    int prog_last = -1;
    auto prog_call = [&](int pct) {
        if (progress && pct != prog_last) { prog_last = pct; progress(pct, progress_user); }
    };
    auto prog_mark = [&](long long its_done, cudaStream_t st) {
        if (!progress) return;
        cudaEvent_t e;
        if (cudaEventCreateWithFlags(&e, cudaEventDisableTiming) != cudaSuccess) { cudaGetLastError(); return; }
        cudaEventRecord(e, st);
        prog_ev.emplace_back(e, (int) std::min<long long>(99, its_done * 100 / std::max<long long>(1, iterations)));
    };
    auto prog_drain = [&]() {
        for (auto &pe : prog_ev) {
            cudaError_t q;
            while ((q = cudaEventQuery(pe.first)) == cudaErrorNotReady) std::this_thread::sleep_for(std::chrono::microseconds(200));
            if (q != cudaSuccess) cudaGetLastError();
            else prog_call(pe.second);
            cudaEventDestroy(pe.first);
        }
        prog_ev.clear();
    };
    prog_call(0);
    auto run_iteration = [&](long long it) {
        for (int g = 0; g < G; ++g) {
            Group &q = grp[g];
            unsigned *list = q.lists[q.cur], *n_list = q.n_list[q.cur];
            mark();
#define RTB_LAUNCH_T(COUNT, MINB, MODE) k_wf_trace<COUNT, MINB, MODE><<<q.trace_grid, 128, 0, q.st>>>(sc, w, list, n_list, q.queue_head, p.d_counters, pol, rays_per_lane, tail_steps)
#define RTB_LAUNCH_TM(MODE)                                          \
            if (p.count_tests) RTB_LAUNCH_T(true, 10, MODE);         \
            else if (bps <= 9) RTB_LAUNCH_T(false, 9, MODE);         \
            else RTB_LAUNCH_T(false, 10, MODE)
            if (mode == 3) { RTB_LAUNCH_TM(3); } else if (mode == 5) { RTB_LAUNCH_TM(5); } else { RTB_LAUNCH_TM(4); }
#undef RTB_LAUNCH_TM
#undef RTB_LAUNCH_T
            mark();
            if (p.count_tests) k_wf_shade<true, 5><<<q.shade_grid, kShadeBlock, 0, q.st>>>(sc, p, w, list, n_list, q.queue_head);
            else if (shade_bps <= 5) k_wf_shade<false, 5><<<q.shade_grid, kShadeBlock, 0, q.st>>>(sc, p, w, list, n_list, q.queue_head);
            else if (shade_bps <= 6) k_wf_shade<false, 6><<<q.shade_grid, kShadeBlock, 0, q.st>>>(sc, p, w, list, n_list, q.queue_head);
            else k_wf_shade<false, 8><<<q.shade_grid, kShadeBlock, 0, q.st>>>(sc, p, w, list, n_list, q.queue_head);
            nl += 2;
            if ((it + 1) % period == 0 && it + 1 < iterations) compact(q);
        }
    };
    // The groups' iterations are many short launches (1/8 of a 1080p frame in 8 groups: 20 k launches per 64 spp, and
    // the host thread tops out at ~3 us per launch: profiles/r01k_groups_sweep.txt).  So 32 iterations -- two compaction
    // periods, after which every group's list pointers are back where they started -- are captured ONCE into a CUDA
    // graph (the groups' streams fork from and join the caller's stream inside the capture) and replayed; what is left
    // over is launched directly.  Every kernel reads its live list length from device memory, so a replay is valid.
    long long it0 = 0;
    static const bool graphs_on = [] { const char *e = std::getenv("RTB_WF_GRAPH"); return !(e && std::atoi(e) == 0); }();
    const long long kBlockIts = 2 * period;
    const long long n_graph_blocks = (G > 1 && !stage_events && graphs_on && iterations > 2 * kBlockIts) ? (iterations - 1) / kBlockIts : 0;
    if (n_graph_blocks > 0) {
        // everything queued on the groups' streams so far (init + first compaction) must precede the first replay
        for (int g = 0; g < G; ++g) {
            cudaEventRecord(aux->join[g], grp[g].st);
            cudaStreamWaitEvent(stream, aux->join[g], 0);
        }
        // the block depends on nothing but the launch parameters: keep it across frames
        std::vector<unsigned char> key;
        auto add = [&](const void *ptr, size_t bytes) { key.insert(key.end(), (const unsigned char *) ptr, (const unsigned char *) ptr + bytes); };
        add(&sc, sizeof sc); add(&p, sizeof p); add(&state, sizeof state); add(&G, sizeof G); add(&mode, sizeof mode);
        add(&bps, sizeof bps); add(&tail_steps, sizeof tail_steps); add(&shade_bps, sizeof shade_bps); add(&rays_per_lane, sizeof rays_per_lane); add(&pol, sizeof pol);
        bool ok = true;
        if (!aux->graph_exec || key != aux->graph_key) {
            if (aux->graph_exec) { cudaStreamSynchronize(stream); cudaGraphExecDestroy(aux->graph_exec); aux->graph_exec = nullptr; }
            if (aux->graph) { cudaGraphDestroy(aux->graph); aux->graph = nullptr; }
            const unsigned long long nl_before = nl;
            int cur_before[kWfMaxGroups];
            for (int g = 0; g < G; ++g) cur_before[g] = grp[g].cur;
            ok = cudaStreamBeginCapture(stream, cudaStreamCaptureModeThreadLocal) == cudaSuccess;
            if (ok) {
                cudaEventRecord(aux->fork, stream);
                for (int g = 0; g < G; ++g) cudaStreamWaitEvent(grp[g].st, aux->fork, 0);
                for (long long it = 0; it < kBlockIts; ++it) run_iteration(it);   // (it + 1 < iterations holds: see n_graph_blocks)
                for (int g = 0; g < G; ++g) {
                    cudaEventRecord(aux->join[g], grp[g].st);
                    cudaStreamWaitEvent(stream, aux->join[g], 0);
                }
                ok = cudaStreamEndCapture(stream, &aux->graph) == cudaSuccess && aux->graph != nullptr;
            }
            if (ok) ok = cudaGraphInstantiate(&aux->graph_exec, aux->graph, 0) == cudaSuccess;
            aux->graph_launches = nl - nl_before;
            nl = nl_before;
            for (int g = 0; g < G; ++g) grp[g].cur = cur_before[g];   // (a whole block leaves them unchanged anyway)
            if (ok) {
                aux->graph_key.swap(key);
            } else {
                cudaGetLastError();   // capture / instantiation unavailable: everything is launched directly below
                if (aux->graph_exec) { cudaGraphExecDestroy(aux->graph_exec); aux->graph_exec = nullptr; }
                if (aux->graph) { cudaGraphDestroy(aux->graph); aux->graph = nullptr; }
                aux->graph_key.clear();
            }
        }
        if (ok) {
            long long done = 0;
            const long long blocks_per_mark = std::max<long long>(1, n_graph_blocks / 100);
            for (; done < n_graph_blocks; ++done) {
                if (cudaGraphLaunch(aux->graph_exec, stream) != cudaSuccess) { cudaGetLastError(); break; }
                if ((done + 1) % blocks_per_mark == 0) prog_mark((done + 1) * kBlockIts, stream);
            }
            nl += aux->graph_launches * (unsigned long long) done;
            it0 = done * kBlockIts;
        }
        // the directly launched rest follows the replays
        cudaEventRecord(aux->fork, stream);
        for (int g = 0; g < G; ++g) cudaStreamWaitEvent(grp[g].st, aux->fork, 0);
    }
    for (long long it = it0; it < iterations; ++it) {
        run_iteration(it);
        if ((it + 1) % prog_stride == 0) prog_mark(it + 1, grp[0].st);
    }
    prog_drain();
    auto join = [&]() {
        if (G <= 1) return;
        for (int g = 0; g < G; ++g) {
            cudaEventRecord(aux->join[g], grp[g].st);
            cudaStreamWaitEvent(stream, aux->join[g], 0);
        }
    };
    if (tail_steps > 0) {
        // Suspended rays cost their slots extra iterations: compact (that counts the live slots), read the counts back,
        // and go on in rounds of 16 until every list is empty.  The margin above makes one read-back the normal case.
        long long it = iterations;
        status = 2;   // until a read-back says every list is empty
        for (int round = 0; round < 4096; ++round) {
            unsigned live[kWfMaxGroups];
            for (int g = 0; g < G; ++g) {
                compact(grp[g]);
                cudaMemcpyAsync(&live[g], grp[g].n_list[grp[g].cur], sizeof(unsigned), cudaMemcpyDeviceToHost, grp[g].st);
            }
            join();
            if (cudaStreamSynchronize(stream) != cudaSuccess) { status = 1; break; }
            unsigned total = 0;
            for (int g = 0; g < G; ++g) total += live[g];
            if (total == 0) { status = 0; break; }
            if (G > 1) {
                cudaEventRecord(aux->fork, stream);
                for (int g = 0; g < G; ++g) cudaStreamWaitEvent(grp[g].st, aux->fork, 0);
            }
            for (int k = 0; k < 16; ++k) run_iteration(it++);
        }
    }
    mark();
    join();
    if (p.substreams > 1) {   // every stream of every pixel has finished: fold them into the pixels' sums
        const size_t npix = (size_t) p.local_rows * (size_t) p.width;
        k_wf_fold<<<(unsigned) ((npix + 255) / 256), 256, 0, stream>>>(p, w);
        ++nl;
    }
    if (progress) {   // (with a progress line the call returns when the frame is complete)
        cudaStreamSynchronize(stream);
        prog_call(100);
    }
    if (stage_events) {   // events beyond ev_i were created by an earlier, longer frame
        for (size_t i = ev_i; i < stage_events->size(); ++i) cudaEventDestroy((*stage_events)[i]);
        stage_events->resize(ev_i);
    }
    if (launches) *launches += nl;
    return status;
}

}  // namespace rtb
