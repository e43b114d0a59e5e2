// Test tooling: a host stand-in for the NCCL entry points csrc/rtb_dist.cu loads with dlopen (RTB_NCCL_LIB), with NCCL's
// blocking behaviour: a send waits for its receive, a reduce for every rank -- a rank that never shows up leaves the others
// waiting, which is exactly what tests/devsim/dist_emul_check.cpp wants to be able to see.  "Device" buffers are host memory.
#include <condition_variable>
#include <cstdint>
#include <cstring>
#include <map>
#include <mutex>
#include <string>
#include <vector>

namespace {
struct Id { char b[128]; };
struct World {
    int n = 0, joined = 0;
    std::mutex m;
    std::condition_variable cv;
    struct Slot { const void *buf = nullptr; size_t bytes = 0; bool posted = false; };
    std::vector<Slot> mail;                 // [src * n + dst]
    std::vector<const void *> red_in;       // reduce: every rank's input
    int red_arrived = 0, red_gen = 0;
};
std::mutex g_mu;
std::map<std::string, World *> g_worlds;
int g_next_id = 1;
size_t type_bytes(int t) { return t == 7 ? 4 : 1; }   // ncclFloat32 = 7, ncclUint8 = 1
}  // namespace

struct ncclComm { World *w; int rank; };

extern "C" {
int ncclGetVersion(int *v) { *v = 22809; return 0; }
const char *ncclGetErrorString(int) { return "fake NCCL error"; }
int ncclGetUniqueId(Id *id) {
    std::lock_guard<std::mutex> lk(g_mu);
    std::memset(id, 0, sizeof *id);
    std::snprintf(id->b, sizeof id->b, "fake-%d", g_next_id++);
    return 0;
}
int ncclCommInitRank(ncclComm **out, int n, Id id, int rank) {
    World *w;
    {
        std::lock_guard<std::mutex> lk(g_mu);
        World *&slot = g_worlds[std::string(id.b)];
        if (!slot) { slot = new World(); slot->n = n; slot->mail.resize((size_t) n * n); slot->red_in.resize((size_t) n); }
        w = slot;
    }
    std::unique_lock<std::mutex> lk(w->m);
    ++w->joined;
    w->cv.notify_all();
    w->cv.wait(lk, [&] { return w->joined >= w->n; });   // as ncclCommInitRank: returns when every rank has joined
    *out = new ncclComm{w, rank};
    return 0;
}
int ncclCommDestroy(ncclComm *c) { delete c; return 0; }
int ncclGroupStart() { return 0; }
int ncclGroupEnd() { return 0; }
int ncclSend(const void *buf, size_t count, int type, int peer, ncclComm *c, void *) {
    World *w = c->w;
    std::unique_lock<std::mutex> lk(w->m);
    World::Slot &s = w->mail[(size_t) c->rank * w->n + peer];
    s.buf = buf; s.bytes = count * type_bytes(type); s.posted = true;
    w->cv.notify_all();
    w->cv.wait(lk, [&] { return !s.posted; });           // until the receive has taken it
    return 0;
}
int ncclRecv(void *buf, size_t count, int type, int peer, ncclComm *c, void *) {
    World *w = c->w;
    std::unique_lock<std::mutex> lk(w->m);
    World::Slot &s = w->mail[(size_t) peer * w->n + c->rank];
    w->cv.wait(lk, [&] { return s.posted; });
    if (s.bytes != count * type_bytes(type)) return 3;
    std::memcpy(buf, s.buf, s.bytes);
    s.posted = false;
    w->cv.notify_all();
    return 0;
}
int ncclReduce(const void *in, void *out, size_t count, int type, int op, int root, ncclComm *c, void *) {
    if (type != 7 || op != 0) return 4;
    World *w = c->w;
    std::unique_lock<std::mutex> lk(w->m);
    const int gen = w->red_gen;
    w->red_in[(size_t) c->rank] = in;
    if (++w->red_arrived == w->n) {
        // (summed by whoever arrives last, in rank order, into the root's output: the root is still waiting below)
        w->red_arrived = 0;
        ++w->red_gen;
        w->cv.notify_all();
    } else {
        w->cv.wait(lk, [&] { return w->red_gen != gen; });
    }
    if (c->rank == root) {
        float *o = (float *) out;
        for (size_t i = 0; i < count; ++i) {
            float s = 0.f;
            for (int r = 0; r < w->n; ++r) s += ((const float *) w->red_in[(size_t) r])[i];
            o[i] = s;
        }
    }
    // nobody leaves before the root has read every input
    const int gen2 = w->red_gen;
    if (++w->red_arrived == w->n) { w->red_arrived = 0; ++w->red_gen; w->cv.notify_all(); }
    else w->cv.wait(lk, [&] { return w->red_gen != gen2; });
    return 0;
}
}
