// Meeting point of the ranks of one process before a collective (csrc/rtb_dist.cu: rtb_render_multi).  Plain C++ so that
// tests/devsim/multisync_check.cpp can drive it without a GPU.
#pragma once
#include <condition_variable>
#include <mutex>

// The ranks of ONE process (rtb_render_multi: a thread per device) meet here between "my part of the frame is rendered" and
// the collective: a rank that failed before the collective would otherwise leave the others waiting in NCCL for ever.  Every
// rank arrives with its status; all of them leave with the first failure (or 0) and enter the collective together or not
// at all.  (Ranks in different processes have no such meeting point: a failed rank there ends its process.)
struct MultiSync {
    std::mutex m;
    std::condition_variable cv;
    int n = 0, arrived = 0, generation = 0, failed_rc = 0;
    bool met[64] = {};            // rank r has been here (a rank that returns early is brought here by rtb_render_multi)
    int arrive(int rank, int rc) {
        std::unique_lock<std::mutex> lk(m);
        met[rank] = true;
        if (rc != 0 && failed_rc == 0) failed_rc = rc;
        const int gen = generation;
        if (++arrived == n) {
            arrived = 0;
            ++generation;
            cv.notify_all();
        } else {
            cv.wait(lk, [&] { return generation != gen; });
        }
        return failed_rc;
    }
};

