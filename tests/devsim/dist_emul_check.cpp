// Test tooling: the multi-GPU layer of librtb ITSELF (csrc/rtb_dist.cu, unchanged: rtb_render_multi, rtb_render_dist, the rank
// threads' meeting point, the deferred frees) compiled for the host and run with one thread per "device" against a host stand-in
// for NCCL that blocks as NCCL does (emul/fake_nccl.cpp) and a stub for the per-rank render.  What is checked:
//   * tile split and sample split frames arrive complete on rank 0 (rows in place, sums of every rank);
//   * a rank whose render fails makes rtb_render_multi return an error -- the others do not wait in the collective for ever --
//     and the kept communicators render the next frame;
//   * no rank thread ever calls cudaFree, however full the block cache is when a scratch block grows (the 8-GPU hang of r02l).
#define RTB_BUILD_EMULATE 1
#include "../../raytracing-hw_b200/csrc/rtb_block_cache.h"
#include "dist_emul_gen.inc"

#include <atomic>
#include <csignal>
#include <cstdio>
#include <unistd.h>

struct rtb_scene {
    int device = 0;
    void *scratch[3] = {nullptr, nullptr, nullptr};
    size_t scratch_bytes[3] = {0, 0, 0};
    unsigned long long launches = 0;
    int fail = 0;
};

namespace {
thread_local std::string t_error;
rtb::BlockCache g_blocks;
std::atomic<int> g_frees_on_rank_threads{0};
std::thread::id g_main_thread;
void free_hook() { if (std::this_thread::get_id() != g_main_thread) ++g_frees_on_rank_threads; }   // (rtb_render_multi's threads are the only others)
}  // namespace

namespace rtb {
void set_error(const std::string &e) { t_error = e; }
void blocks_defer_frees(bool on) { BlockCache::defer_flag() = on && !std::getenv("DIST_EMUL_NO_DEFER"); }   // (the variable: what the hang looked like)
void blocks_flush_deferred() { g_blocks.flush_deferred(); }
cudaStream_t scene_stream(rtb_scene *) { return nullptr; }
// as csrc/rtb_api.cu: a scratch block that has to grow goes back to the cache, a new one comes from it
int scene_scratch(rtb_scene *sc, int which, size_t bytes, void **out) {
    if (sc->scratch_bytes[which] < bytes) {
        if (sc->scratch[which]) g_blocks.release(sc->device, sc->scratch[which], sc->scratch_bytes[which]);
        sc->scratch[which] = nullptr;
        sc->scratch_bytes[which] = 0;
        size_t got = 0;
        if (g_blocks.acquire(sc->device, bytes, &sc->scratch[which], &got) != cudaSuccess) { set_error("out of memory"); return RTB_ERR_CUDA; }
        sc->scratch_bytes[which] = got;
    }
    *out = sc->scratch[which];
    return RTB_OK;
}
void launch_resolve(const float *accum, size_t n_pixels, int, uint8_t *rgb8, cudaStream_t) {
    for (size_t i = 0; i < n_pixels * 3; ++i) rgb8[i] = (uint8_t) accum[i];
}
}  // namespace rtb

extern "C" {
const char *rtb_last_error(void) { return t_error.c_str(); }
int rtb_device_count(void) { return 8; }
void rtb_release_cached_memory(void) { g_blocks.flush_deferred(); g_blocks.trim(-1); }   // as csrc/rtb_api.cu
int rtb_scene_device(const rtb_scene *s) { return s->device; }
unsigned long long rtb_scene_launches(const rtb_scene *s) { return s->launches; }
int rtb_render_local_rows(const rtb_render_params *p) {   // as csrc/rtb_api.cu
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
static int global_row(const rtb_render_params *p, int lr) {
    if (p->interleave_count <= 1) return p->row_begin + lr;
    const int R = p->interleave_rows;
    return p->row_begin + ((lr / R) * p->interleave_count + p->interleave_index) * R + lr % R;
}
// the stand-in for a rank's render: row y of the frame is filled with y (RGB8) / every float with substream_begin + 1 (sums)
int rtb_render_device(rtb_scene *scene, const rtb_render_params *p, float *d_accum, uint8_t *d_rgb8, unsigned long long *d_counters, void *) {
    if (scene->fail) { rtb::set_error("injected render failure"); return RTB_ERR_CUDA; }
    ++scene->launches;
    const int rows = rtb_render_local_rows(p);
    if (d_rgb8)
        for (int lr = 0; lr < rows; ++lr) std::memset(d_rgb8 + (size_t) lr * p->width * 3, global_row(p, lr) & 255, (size_t) p->width * 3);
    if (d_accum)
        for (size_t i = 0; i < (size_t) rows * p->width * 3; ++i) d_accum[i] = (float) (p->substream_begin + 1);
    if (d_counters) d_counters[0] += (unsigned long long) rows * p->width;
    return RTB_OK;
}
}

static void on_alarm(int) {
    const char msg[] = "FAILED: a frame did not return within 60 s (ranks waiting in the collective)\n";
    if (write(1, msg, sizeof msg - 1)) {}
    _exit(3);
}

int main(int argc, char **argv) {
    setenv("RTB_NCCL_LIB", argc > 1 ? argv[1] : "./libfakenccl.so", 1);
    setenv("RTB_BLOCK_CACHE_MB", "128", 1);   // (the prefill below overfills it: every later release has to drop something)
    signal(SIGALRM, on_alarm);
    emu_free_hook = &free_hook;
    g_main_thread = std::this_thread::get_id();
    int bad = 0;
    auto expect = [&](bool ok, const char *what) { std::printf("%s  %s\n", ok ? "ok" : "FAILED", what); if (!ok) bad = 1; };
    const int N = 8, W = 322, H = 181;
    std::vector<rtb_scene> sc((size_t) N);
    std::vector<rtb_scene *> ptr;
    for (int d = 0; d < N; ++d) { sc[(size_t) d].device = d; ptr.push_back(&sc[(size_t) d]); }
    // a cache that is full of OTHER devices' blocks: every release from here on has to drop something
    auto fill_cache = [&] {
        for (int i = 0; i < 40; ++i) {
            void *p = nullptr;
            size_t got = 0;
            g_blocks.acquire(i % N, (size_t) (6 + i) << 20, &p, &got);   // (growing sizes: none is handed out again)
            g_blocks.release(i % N, p, got);
        }
    };
    fill_cache();
    int freed_before = 0;
    for (int v : emu_frees) freed_before += v;
    rtb_render_params f;
    std::memset(&f, 0, sizeof f);
    f.width = W; f.height = H; f.samples = 6; f.row_end = H; f.pixel0_seed = 1;
    std::vector<uint8_t> img((size_t) W * H * 3);
    rtb_stats st;
    auto rows_ok = [&] { for (int y = 0; y < H; ++y) for (int x = 0; x < W * 3; ++x) if (img[(size_t) y * W * 3 + x] != (uint8_t) (y & 255)) return false; return true; };

    alarm(60);
    int rc = rtb_render_multi(ptr.data(), N, &f, RTB_SPLIT_TILE, img.data(), &st);
    expect(rc == RTB_OK && rows_ok() && st.closest_hit_rays == (uint64_t) W * H, "tile split: every rank's rows in place on rank 0");
    alarm(60);
    std::fill(img.begin(), img.end(), 0);
    f.substreams = N;
    rc = rtb_render_multi(ptr.data(), N, &f, RTB_SPLIT_SAMPLE, img.data(), &st);   // (the scratch blocks grow here: releases into the full cache)
    bool sums = rc == RTB_OK;
    for (uint8_t v : img) sums = sums && v == N * (N + 1) / 2;
    expect(sums, "sample split: the sums of all ranks reduced and resolved on rank 0");
    alarm(60);
    f.substreams = 0;
    sc[3].fail = 1;
    rc = rtb_render_multi(ptr.data(), N, &f, RTB_SPLIT_TILE, img.data(), &st);
    expect(rc != RTB_OK && std::string(rtb_last_error()).find("rank") != std::string::npos, "a failed rank: an error, nobody left in the collective");
    alarm(60);
    f.substreams = N;
    rc = rtb_render_multi(ptr.data(), N, &f, RTB_SPLIT_SAMPLE, img.data(), &st);
    expect(rc != RTB_OK, "the same in the sample split");
    alarm(60);
    sc[3].fail = 0;
    fill_cache();   // (a failed frame hands the cache back to the driver -- on the calling thread: fill it again)
    f.substreams = 0;
    std::fill(img.begin(), img.end(), 0);
    rc = rtb_render_multi(ptr.data(), N, &f, RTB_SPLIT_TILE, img.data(), &st);
    expect(rc == RTB_OK && rows_ok(), "the kept communicators render the next frame");
    // bigger canvases: more growth, more drops
    for (int k = 0; k < 3; ++k) {
        alarm(60);
        f.width = W + 64 * (k + 1); f.height = H + 40 * (k + 1); f.row_end = f.height;
        img.assign((size_t) f.width * f.height * 3, 0);
        f.substreams = k % 2 ? N : 0;
        rc = rtb_render_multi(ptr.data(), N, &f, k % 2 ? RTB_SPLIT_SAMPLE : RTB_SPLIT_TILE, img.data(), &st);
        if (rc != RTB_OK) bad = 1;
    }
    // a NEW device list (communicators have to be created) with one device that cannot be selected: the other ranks must not
    // be left in ncclCommInitRank, which waits for all of them (the alarm ends the test if they are)
    {
        alarm(60);
        std::vector<rtb_scene *> rev(ptr.rbegin(), ptr.rend());
        emu_bad_device = 5;
        f.width = W; f.height = H; f.row_end = H; f.substreams = 0;
        img.assign((size_t) W * H * 3, 0);
        rc = rtb_render_multi(rev.data(), N, &f, RTB_SPLIT_TILE, img.data(), &st);
        expect(rc != RTB_OK && std::string(rtb_last_error()).find("rank") != std::string::npos, "a rank without its device: an error before the communicators are created");
        emu_bad_device = -1;
        rc = rtb_render_multi(rev.data(), N, &f, RTB_SPLIT_TILE, img.data(), &st);
        expect(rc == RTB_OK && rows_ok(), "the same device list renders once the device is back");
        rtb_scene *twice[2] = {ptr[1], ptr[1]};
        rc = rtb_render_multi(twice, 2, &f, RTB_SPLIT_TILE, img.data(), &st);
        expect(rc == RTB_ERR_INVALID, "two ranks on one device are refused");
    }
    alarm(0);
    expect(g_frees_on_rank_threads.load() == 0, "no cudaFree on a thread that holds a rank");
    int freed = 0;
    for (int v : emu_frees) freed += v;
    std::printf("    (frees %d, deferred %zu, cached %zu blocks, %zu MB)\n", freed, g_blocks.deferred.size(), g_blocks.blocks.size(), g_blocks.total >> 20);
    expect(freed > freed_before && g_blocks.deferred.empty(), "what the rank threads dropped was freed after the threads had joined");
    std::printf("%s\n", bad ? "FAILED" : "ok");
    return bad;
}
