// `solution input.gltf width height samples [output.ppm]` -- the reference's command line (src/main.cpp:22-58,
// run.sh:2) in front of librtb.so.  Same argument rules, same stdout lines in the same order, same output file;
// errors abort through an uncaught std::runtime_error exactly like the reference (exit 134).
//
// Scene load + BVH build run on the host (rtb_host_scene_*); the render runs on every visible CUDA device (or the
// first RT_GPUS of them): one host thread per device, rows dealt out in interleaved blocks so that every pixel keeps
// its own sequential sample stream and the image does not depend on the number of devices.
//
// Environment knobs (the CLI itself stays drop-in):  RT_GPUS=n   RT_CULL=reference|global   RT_SEED0=<pixel-0 seed>
//                                                    RT_STATS=1 (print a JSON perf record on stderr)
#include "../../include/rtb.h"

#include <chrono>
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
    rtb_scene_desc desc;
    check(rtb_host_scene_desc(hs, &desc), "desc");

    int ndev = rtb_device_count();
    if (ndev <= 0) throw std::runtime_error("no CUDA device available (this build has no CPU fallback)");
    if (const char *e = std::getenv("RT_GPUS")) ndev = std::max(1, std::min(ndev, std::atoi(e)));
    std::vector<rtb_scene *> scenes((size_t) ndev, nullptr);
    for (int d = 0; d < ndev; ++d) check(rtb_scene_create(&desc, d, &scenes[(size_t) d]), "upload");
    std::cout << "Scene loaded: " << std::setprecision(2) << seconds_since(t_all) << " seconds." << std::endl;

    std::cout << std::setprecision(6) << "Rendering scene." << std::endl;
    std::cout << "Render launched." << std::endl;
    int flags = 0;
    if (const char *e = std::getenv("RT_CULL")) if (std::strcmp(e, "reference") == 0) flags |= RTB_RENDER_REFERENCE_CULL;
    unsigned seed0 = 1;
    if (const char *e = std::getenv("RT_SEED0")) seed0 = (unsigned) std::strtoul(e, nullptr, 10);

    std::vector<uint8_t> image((size_t) width * height * 3);
    std::vector<rtb_stats> stats((size_t) ndev);
    std::vector<std::string> errors((size_t) ndev);
    const auto t_render = Clock::now();
    {
        std::vector<std::thread> workers;
        for (int d = 0; d < ndev; ++d) {
            workers.emplace_back([&, d] {
                rtb_render_params p;
                std::memset(&p, 0, sizeof(p));
                p.width = width; p.height = height; p.samples = samples;
                p.row_begin = 0; p.row_end = height;
                p.pixel0_seed = seed0;
                p.flags = flags;
                p.interleave_count = ndev; p.interleave_index = d; p.interleave_rows = 8;
                const int rows = rtb_render_local_rows(&p);
                std::vector<uint8_t> part((size_t) std::max(rows, 0) * width * 3);
                const int rc = rtb_render(scenes[(size_t) d], &p, nullptr, part.data(), &stats[(size_t) d]);
                if (rc != RTB_OK) { errors[(size_t) d] = rtb_last_error(); return; }
                for (int lr = 0; lr < rows; ++lr)
                    std::memcpy(&image[(size_t) rtb_render_global_row(&p, lr) * width * 3], &part[(size_t) lr * width * 3], (size_t) width * 3);
            });
        }
        for (auto &w : workers) w.join();
    }
    for (const auto &e : errors) if (!e.empty()) throw std::runtime_error("render: " + e);
    const double render_s = seconds_since(t_render);
    progress_done(seconds_since(t_all));   // the reference's timer also includes the load (src/main.cpp:44,54)

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
