// `solution input.gltf width height samples [output.ppm]` -- the reference's command line (src/main.cpp:22-58,
// run.sh:2) in front of librtb.so.  Same argument rules, same stdout lines in the same order, same output file;
// errors abort through an uncaught std::runtime_error exactly like the reference (exit 134).
//
// Scene load + BVH build run on the host (rtb_host_scene_*); the render runs on every visible CUDA device (or the
// first RT_GPUS of them) through rtb_render_multi: one rank per device, the frame combined on device 0 over NCCL /
// NVLink and copied to the host once.  RT_SPLIT=tile (default): rows dealt out in interleaved blocks, every pixel
// keeps its own sequential sample stream and the image does not depend on the number of devices.  RT_SPLIT=sample:
// every device renders the whole canvas with its share of RT_SUBSTREAMS (default: the device count) engine streams per
// pixel, partial sums reduced with ncclReduce -- statistically equivalent, not bit-identical, scales on small canvases.
//
// Environment knobs (the CLI itself stays drop-in):  RT_GPUS=n   RT_CULL=reference|global   RT_SEED0=<pixel-0 seed>
//                                                    RT_SPLIT=tile|sample   RT_SUBSTREAMS=<K>
//                                                    RT_STATS=1 (print a JSON perf record on stderr)
//                                                    RT_PROGRESS=1 (the USE_CALLBACK build's progress line)
//                                                    RT_CHUNK=<spp> RT_CHECKPOINT=<file> (progressive frame, resume)
#include "../../include/rtb.h"

#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <iomanip>
#include <iostream>
#include <stdexcept>
#include <string>
#include <thread>
#include <vector>

namespace {

using Clock = std::chrono::steady_clock;
double seconds_since(Clock::time_point t0) { return std::chrono::duration<double>(Clock::now() - t0).count(); }

void check(int rc, const char *what) {
    if (rc != RTB_OK) throw std::runtime_error(std::string(what) + ": " + rtb_last_error());
}

// int_from_string (src/utils/vector.h:67-72): std::stoi on the text up to the first blank
int int_from_string(const std::string &s) {
    const size_t end_pos = s.find(' ', 0);
    return std::stoi(s.substr(0, end_pos));
}

// ProgressCallback(100, t) (src/main.cpp:8-20)
void progress_done(double elapsed) {
    std::cout << "\r   Render [";
    for (int i = 0; i < 10; i++) std::cout << "*";
    std::cout << "] " << 100 << "% (" << elapsed << " s)" << std::flush;
    std::cout << "\n      " << elapsed << " seconds (" << elapsed * 1000 << " ms) elapsed." << std::endl;
}

// ProgressCallback(progress, t) of the reference's USE_CALLBACK build (src/main.cpp:8-20), RT_PROGRESS=1: the timer
// restarts at 0 %, the closing line is printed at 100 %
Clock::time_point g_progress_t0;
int progress_line(int progress, void *) {
    if (progress == 0) g_progress_t0 = Clock::now();
    const double el = seconds_since(g_progress_t0);
    std::cout << "\r   Render [";
    for (int i = 0; i < 10; i++) std::cout << (progress / ((i + 1) * 10) ? "*" : " ");
    std::cout << "] " << progress << "% (" << el << " s)" << std::flush;
    if (progress == 100) std::cout << "\n      " << el << " seconds (" << el * 1000 << " ms) elapsed." << std::endl;
    return 1;
}

// RT_CHECKPOINT=<file>: the frame's per-pixel state (rtb_render_resume) in global row order behind this header
struct CheckpointHeader {
    char magic[8];
    int32_t width, height, samples_done;
    uint32_t seed0;
};
const char kCheckpointMagic[8] = {'R', 'T', 'B', 'C', 'K', 'P', 'T', '1'};

bool read_checkpoint(const std::string &path, int width, int height, unsigned seed0, int samples, std::vector<uint8_t> &state, int &done) {
    FILE *f = std::fopen(path.c_str(), "rb");
    if (!f) return false;
    CheckpointHeader h;
    bool ok = std::fread(&h, sizeof h, 1, f) == 1 && std::memcmp(h.magic, kCheckpointMagic, 8) == 0 && h.width == width &&
              h.height == height && h.seed0 == seed0 && h.samples_done > 0 && h.samples_done < samples;
    if (ok) ok = std::fread(state.data(), 1, state.size(), f) == state.size();
    std::fclose(f);
    if (ok) done = h.samples_done;
    return ok;
}

void write_checkpoint(const std::string &path, int width, int height, unsigned seed0, int done, const std::vector<uint8_t> &state) {
    const std::string tmp = path + ".tmp";
    FILE *f = std::fopen(tmp.c_str(), "wb");
    if (!f) throw std::runtime_error("cannot write " + tmp);
    CheckpointHeader h;
    std::memcpy(h.magic, kCheckpointMagic, 8);
    h.width = width; h.height = height; h.samples_done = done; h.seed0 = seed0;
    const bool ok = std::fwrite(&h, sizeof h, 1, f) == 1 && std::fwrite(state.data(), 1, state.size(), f) == state.size();
    std::fclose(f);
    if (!ok || std::rename(tmp.c_str(), path.c_str()) != 0) throw std::runtime_error("cannot write " + path);
}

}  // namespace

int main(int argc, char *argv[]) {
    if (argc < 5 || argc > 6) throw std::runtime_error("Invalid arguments - " + std::to_string(argc) + " (expected: 5)");
    const std::string input = argv[1];
    const int width = int_from_string(argv[2]);
    const int height = int_from_string(argv[3]);
    const int samples = int_from_string(argv[4]);
    std::string output = "output.ppm";
    if (argc == 6) output = argv[5];

    std::cout << "Loading scene." << std::endl;
    const auto t_all = Clock::now();
    rtb_host_scene *hs = nullptr;
    check(rtb_host_scene_load_gltf(input.c_str(), width, height, samples, &hs), "load");
    {   // Scene::Scene prints the BVH statistics (src/core/scene.cpp:233-246)
        uint64_t n_nodes = 0, counts[5] = {0, 0, 0, 0, 0}, max_leaf = 0;
        rtb_host_scene_bvh_stats(hs, &n_nodes, counts, &max_leaf);
        std::cout << n_nodes << " nodes in BVH." << std::endl;
        std::cout << "Node object count:" << std::endl;
        for (int i = 0; i < 5; ++i) std::cout << "\t" << i << ": " << counts[i] << std::endl;
        std::cout << "At most " << max_leaf << " objects in single node." << std::endl;
    }
    // the triangles are assembled on the GPU from the file's own views (rtb_scene_create_raw); RT_ASSEMBLE=host: the
    // host assembles them as the reference does while parsing (scene_parser.cpp:280-345) and uploads the result
    rtb_scene_desc desc;
    rtb_raw_geometry raw;
    const char *asm_env = std::getenv("RT_ASSEMBLE");
    const bool host_assembly = asm_env && std::strcmp(asm_env, "host") == 0;
    if (host_assembly) check(rtb_host_scene_desc(hs, &desc), "desc");
    else check(rtb_host_scene_raw_geometry(hs, &desc, &raw), "desc");

    int ndev = rtb_device_count();
    if (ndev <= 0) throw std::runtime_error("no CUDA device available (this build has no CPU fallback)");
    if (const char *e = std::getenv("RT_GPUS")) ndev = std::max(1, std::min(ndev, std::atoi(e)));
    std::vector<rtb_scene *> scenes((size_t) ndev, nullptr);
    for (int d = 0; d < ndev; ++d)
        check(host_assembly ? rtb_scene_create(&desc, d, &scenes[(size_t) d]) : rtb_scene_create_raw(&desc, &raw, d, &scenes[(size_t) d]), "upload");
    std::cout << "Scene loaded: " << std::setprecision(2) << seconds_since(t_all) << " seconds." << std::endl;

    std::cout << std::setprecision(6) << "Rendering scene." << std::endl;
    std::cout << "Render launched." << std::endl;
    int flags = 0;
    if (const char *e = std::getenv("RT_CULL")) if (std::strcmp(e, "reference") == 0) flags |= RTB_RENDER_REFERENCE_CULL;
    unsigned seed0 = 1;
    if (const char *e = std::getenv("RT_SEED0")) seed0 = (unsigned) std::strtoul(e, nullptr, 10);

    // Progressive rendering (off by default: one call per device, no state).  RT_CHUNK=<n>: the frame is rendered n
    // samples at a time and the output image is rewritten after every chunk (each is the finished image at that sample
    // count); RT_CHECKPOINT=<file>: the per-pixel state is stored after every chunk and a run that finds a matching
    // file carries on from it.  The finished frame has the bits of the one-call frame either way.
    const char *ckpt_env = std::getenv("RT_CHECKPOINT");
    const std::string ckpt = ckpt_env ? ckpt_env : "";
    int chunk = 0;
    if (const char *e = std::getenv("RT_CHUNK")) chunk = std::max(0, std::atoi(e));
    const bool progressive = (!ckpt.empty() || chunk > 0) && !(flags & RTB_RENDER_REFERENCE_CULL);
    const bool progress_on = std::getenv("RT_PROGRESS") != nullptr && std::atoi(std::getenv("RT_PROGRESS")) != 0;
    if (progress_on) check(rtb_scene_set_progress(scenes[0], progress_line, nullptr), "progress");

    std::vector<uint8_t> image((size_t) width * height * 3);
    std::vector<uint8_t> frame_state(progressive ? (size_t) width * height * RTB_PIXEL_STATE_BYTES : 0);
    std::vector<rtb_stats> stats((size_t) ndev);
    std::vector<std::string> errors((size_t) ndev);
    int done = 0;
    if (progressive && !ckpt.empty() && read_checkpoint(ckpt, width, height, seed0, samples, frame_state, done))
        std::cerr << "resuming from " << ckpt << ": " << done << " of " << samples << " samples per pixel done" << std::endl;
    const auto t_render = Clock::now();
    if (!progressive) {
        // the whole frame in one collective call: per-device render, NCCL combine on device 0, one device->host copy
        rtb_render_params p;
        std::memset(&p, 0, sizeof(p));
        p.width = width; p.height = height; p.samples = samples;
        p.row_begin = 0; p.row_end = height;
        p.pixel0_seed = seed0;
        p.flags = flags;
        p.interleave_rows = 8;
        int split = RTB_SPLIT_TILE;
        if (const char *e = std::getenv("RT_SPLIT")) if (std::strcmp(e, "sample") == 0) split = RTB_SPLIT_SAMPLE;
        if (const char *e = std::getenv("RT_SUBSTREAMS")) p.substreams = std::max(0, std::atoi(e));
        rtb_stats all;
        check(rtb_render_multi(scenes.data(), ndev, &p, split, image.data(), &all), "render");
        stats[0] = all;
        done = samples;
    }
    while (done < samples) {
        const int next = (progressive && chunk > 0) ? std::min(samples, done + chunk) : samples;
        std::vector<std::thread> workers;
        for (int d = 0; d < ndev; ++d) {
            workers.emplace_back([&, d] {
                rtb_render_params p;
                std::memset(&p, 0, sizeof(p));
                p.width = width; p.height = height; p.samples = next;
                p.sample_begin = done;
                p.row_begin = 0; p.row_end = height;
                p.pixel0_seed = seed0;
                p.flags = flags;
                p.interleave_count = ndev; p.interleave_index = d; p.interleave_rows = 8;
                const int rows = rtb_render_local_rows(&p);
                std::vector<uint8_t> part((size_t) std::max(rows, 0) * width * 3);
                const size_t srow = (size_t) width * RTB_PIXEL_STATE_BYTES;
                std::vector<uint8_t> st(progressive ? (size_t) std::max(rows, 0) * srow : 0);
                if (progressive && done > 0)
                    for (int lr = 0; lr < rows; ++lr) std::memcpy(&st[(size_t) lr * srow], &frame_state[(size_t) rtb_render_global_row(&p, lr) * srow], srow);
                rtb_stats one;
                const int rc = progressive ? rtb_render_resume(scenes[(size_t) d], &p, st.data(), nullptr, part.data(), &one)
                                           : rtb_render(scenes[(size_t) d], &p, nullptr, part.data(), &one);
                if (rc != RTB_OK) { errors[(size_t) d] = rtb_last_error(); return; }
                rtb_stats &acc = stats[(size_t) d];
                acc.closest_hit_rays += one.closest_hit_rays; acc.light_pdf_queries += one.light_pdf_queries;
                acc.kernel_ms += one.kernel_ms; acc.launches += one.launches;
                for (int lr = 0; lr < rows; ++lr) {
                    const size_t gr = (size_t) rtb_render_global_row(&p, lr);
                    std::memcpy(&image[gr * width * 3], &part[(size_t) lr * width * 3], (size_t) width * 3);
                    if (progressive) std::memcpy(&frame_state[gr * srow], &st[(size_t) lr * srow], srow);
                }
            });
        }
        for (auto &w : workers) w.join();
        for (const auto &e : errors) if (!e.empty()) throw std::runtime_error("render: " + e);
        done = next;
        if (progressive && done < samples) {
            check(rtb_write_ppm(output.c_str(), width, height, image.data()), "write");   // the image so far
            if (!ckpt.empty()) write_checkpoint(ckpt, width, height, seed0, done, frame_state);
        }
    }
    if (progressive && !ckpt.empty()) std::remove(ckpt.c_str());   // the frame is complete
    const double render_s = seconds_since(t_render);
    if (!progress_on) progress_done(seconds_since(t_all));   // the reference's timer also includes the load (src/main.cpp:44,54)

    check(rtb_write_ppm(output.c_str(), width, height, image.data()), "write");
    std::cout << "Frame drawn into " << output << std::endl;

    if (std::getenv("RT_STATS")) {
        uint64_t rays = 0, lq = 0;
        double kernel_ms = 0;
        for (const auto &s : stats) { rays += s.closest_hit_rays; lq += s.light_pdf_queries; kernel_ms = std::max(kernel_ms, s.kernel_ms); }
        std::cerr << "{\"gpus\": " << ndev << ", \"render_s\": " << render_s << ", \"kernel_ms_max\": " << kernel_ms
                  << ", \"closest_hit_rays\": " << rays << ", \"light_pdf_queries\": " << lq
                  << ", \"mrays_per_s\": " << (double) rays / render_s * 1e-6
                  << ", \"mpix_spp_per_s\": " << (double) width * height * samples / render_s * 1e-6
                  << ", \"host_build_s\": " << rtb_host_scene_build_seconds(hs) << "}" << std::endl;
    }
    for (auto *s : scenes) rtb_scene_destroy(s);
    rtb_host_scene_destroy(hs);
    return 0;
}
