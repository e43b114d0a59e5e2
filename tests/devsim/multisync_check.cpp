// Test tooling: the meeting point of rtb_render_multi's rank threads (csrc/rtb_multisync.h) under load -- every rank leaves
// with the first failure of the frame (or 0), nobody waits for a rank that failed early, frames do not leak into each other.
#include <atomic>
#include <cstdio>
#include <random>
#include <thread>
#include <vector>
#include "../../raytracing-hw_b200/csrc/rtb_multisync.h"

int main() {
    int bad = 0;
    for (int frame = 0; frame < 400 && !bad; ++frame) {
        const int n = 2 + frame % 7;
        const int failing = (frame % 3 == 0) ? frame % n : -1;        // one rank fails in every third frame
        const bool early = frame % 2 == 0;                            // ... before it reaches the meeting point (late arrival by the caller)
        MultiSync sync;
        sync.n = n;
        std::vector<int> got((size_t) n, -12345);
        std::vector<std::thread> th;
        for (int r = 0; r < n; ++r)
            th.emplace_back([&, r] {
                std::mt19937 g((unsigned) (frame * 64 + r));
                if (g() % 3 == 0) std::this_thread::sleep_for(std::chrono::microseconds(g() % 300));
                int rc = r == failing ? 7 : 0;
                if (r == failing && early) {
                    // what rtb_render_multi does for a rank whose call returned before the meeting point
                    if (!sync.met[r]) got[(size_t) r] = sync.arrive(r, rc);
                } else {
                    got[(size_t) r] = sync.arrive(r, rc);
                }
            });
        for (auto &t : th) t.join();
        for (int r = 0; r < n; ++r)
            if (got[(size_t) r] != (failing >= 0 ? 7 : 0) || !sync.met[r]) { std::printf("frame %d rank %d: got %d\n", frame, r, got[(size_t) r]); bad = 1; }
    }
    std::printf("%s\n", bad ? "FAILED" : "ok");
    return bad;
}
