// Test tooling: the GPU builder's sort (csrc/rtb_build_steps.cuh: breadth-first replay of libstdc++'s introsort) driven by
// plain loops on the host and compared with std::sort / std::__partial_sort themselves, ties included.
// The loops below mirror the kernels of csrc/rtb_build.cu one to one (same passes, same formulas).
#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <random>
#include <vector>
#include "../../raytracing-hw_b200/csrc/rtb_build_steps.cuh"

using namespace rtb::gbuild;

struct KeyIdx {
    float key;
    uint32_t idx;
};

// one std::sort per segment [seg[s], seg[s + 1]) of (key, idx), all segments at once, round by round
static void sort_rounds(std::vector<float> &key, std::vector<uint32_t> &idx, const std::vector<uint32_t> &seg, int depth_override, long *rounds_out, long *heaps_out) {
    const uint32_t n = (uint32_t) key.size();
    std::vector<SortRange> ranges;
    std::vector<uint32_t> rid(n, kNoRange);
    for (size_t s = 0; s + 1 < seg.size(); ++s) {
        const uint32_t lo = seg[s], hi = seg[s + 1];
        if (hi <= lo) continue;
        const int depth = depth_override >= 0 ? depth_override : 2 * lg_floor(hi - lo);
        const uint32_t id = (uint32_t) ranges.size();
        ranges.push_back({lo, hi, depth, range_state(lo, hi, depth)});
        for (uint32_t i = lo; i < hi; ++i) rid[i] = id;
    }
    long rounds = 0;
    std::vector<unsigned long long> flags(n), scan(n);
    std::vector<uint32_t> posL(n), posR(n);
    std::vector<SortRound> rr;
    while (true) {
        size_t active = 0;
        for (auto &r : ranges) active += r.state == kRangeActive;
        if (!active) break;
        ++rounds;
        const size_t nr = ranges.size();
        rr.assign(nr, SortRound{0.f, 0, kNoRange, 0});
        // K1: pivots
        for (size_t r = 0; r < nr; ++r)
            if (ranges[r].state == kRangeActive) {
                median_to_first(key.data(), idx.data(), ranges[r].lo, ranges[r].hi);
                rr[r].pivot = key[ranges[r].lo];
            }
        // K2: stop flags (low word: left cursor, high word: right cursor)
        for (uint32_t i = 0; i < n; ++i) {
            unsigned long long f = 0;
            const uint32_t r = rid[i];
            if (r != kNoRange && ranges[r].state == kRangeActive && i > ranges[r].lo) {
                if (stops_left(key[i], rr[r].pivot)) f |= 1ull;
                if (stops_right(key[i], rr[r].pivot)) f |= 1ull << 32;
            }
            flags[i] = f;
        }
        // inclusive prefix sum of both words at once
        unsigned long long acc = 0;
        for (uint32_t i = 0; i < n; ++i) { acc += flags[i]; scan[i] = acc; }
        // K3: positions by rank
        for (uint32_t i = 0; i < n; ++i) {
            const unsigned long long f = flags[i];
            if (!f) continue;
            const SortRange &g = ranges[rid[i]];
            const unsigned long long base = scan[g.lo], end = scan[g.hi - 1];
            if (f & 1ull) posL[g.lo + 1 + ((uint32_t) scan[i] - (uint32_t) base - 1u)] = i;
            if (f >> 32) posR[g.lo + 1 + ((uint32_t) (end >> 32) - (uint32_t) (scan[i] >> 32))] = i;   // rank from the right, 0-based
        }
        // K4: exchanges and the cut
        for (uint32_t j = 0; j < n; ++j) {
            const uint32_t r = rid[j];
            if (r == kNoRange || ranges[r].state != kRangeActive || j <= ranges[r].lo) continue;
            const SortRange &g = ranges[r];
            const uint32_t k = j - (g.lo + 1);
            const uint32_t nL = (uint32_t) scan[g.hi - 1] - (uint32_t) scan[g.lo];
            const uint32_t nR = (uint32_t) (scan[g.hi - 1] >> 32) - (uint32_t) (scan[g.lo] >> 32);
            auto valid = [&](uint32_t q) { return q < nL && q < nR && posL[g.lo + 1 + q] < posR[g.lo + 1 + q]; };
            const bool v = valid(k);
            if (k == 0 && !v) rr[r].cut = posL[g.lo + 1];
            if (v) {
                const uint32_t a = posL[j], b = posR[j];
                if (!valid(k + 1)) rr[r].cut = std::min(k + 1 < nL ? posL[j + 1] : kInf, b);
                swap_at(key.data(), idx.data(), a, b);
            }
        }
        // K5: the two parts
        for (size_t r = 0; r < nr; ++r)
            if (ranges[r].state == kRangeActive) {
                SortRange &g = ranges[r];
                const uint32_t cut = rr[r].cut, hi = g.hi;
                const int d = g.depth - 1;
                rr[r].right_id = (uint32_t) ranges.size();
                ranges.push_back({cut, hi, d, range_state(cut, hi, d)});
                SortRange &g2 = ranges[r];
                g2.hi = cut;
                g2.depth = d;
                g2.state = range_state(g2.lo, cut, d);
            }
        // K6: range of every element
        for (uint32_t i = 0; i < n; ++i) {
            const uint32_t r = rid[i];
            if (r != kNoRange && r < nr && rr[r].right_id != kNoRange && i >= rr[r].cut) rid[i] = rr[r].right_id;
        }
    }
    long heaps = 0;
    for (auto &r : ranges) {
        if (r.state == kRangeHeap) { partial_sort_exact(key.data(), idx.data(), r.lo, r.hi); ++heaps; }
        else if (r.state == kRangeSmall) insertion_sort_block(key.data(), idx.data(), r.lo, r.hi);
    }
    if (rounds_out) *rounds_out = rounds;
    if (heaps_out) *heaps_out = heaps;
}

// std::sort with a depth limit of our choosing (what bits/stl_algo.h does, through its own internals)
static void std_sort_depth(KeyIdx *first, KeyIdx *last, long depth) {
    auto comp = __gnu_cxx::__ops::__iter_comp_iter([](const KeyIdx &a, const KeyIdx &b) { return a.key < b.key; });
    if (first == last) return;
    std::__introsort_loop(first, last, depth, comp);
    std::__final_insertion_sort(first, last, comp);
}

static int check(const std::vector<float> &keys, const std::vector<uint32_t> &seg, int depth_override, const char *what) {
    const size_t n = keys.size();
    std::vector<KeyIdx> ref(n);
    for (size_t i = 0; i < n; ++i) ref[i] = {keys[i], (uint32_t) i};
    for (size_t s = 0; s + 1 < seg.size(); ++s) {
        if (depth_override >= 0) std_sort_depth(ref.data() + seg[s], ref.data() + seg[s + 1], depth_override);
        else std::sort(ref.begin() + seg[s], ref.begin() + seg[s + 1], [](const KeyIdx &a, const KeyIdx &b) { return a.key < b.key; });
    }
    std::vector<float> k = keys;
    std::vector<uint32_t> idx(n);
    for (size_t i = 0; i < n; ++i) idx[i] = (uint32_t) i;
    long rounds = 0, heaps = 0;
    sort_rounds(k, idx, seg, depth_override, &rounds, &heaps);
    for (size_t i = 0; i < n; ++i)
        if (idx[i] != ref[i].idx || std::memcmp(&k[i], &ref[i].key, 4) != 0) {
            std::printf("MISMATCH %s: n %zu, element %zu: got idx %u want %u (rounds %ld, heaps %ld)\n", what, n, i, idx[i], ref[i].idx, rounds, heaps);
            return 1;
        }
    return 0;
}

int main(int argc, char **argv) {
    const int reps = argc > 1 ? std::atoi(argv[1]) : 40;
    std::mt19937 g(12345);
    int bad = 0;
    long cases = 0;
    auto sizes = {1u, 2u, 3u, 15u, 16u, 17u, 18u, 31u, 33u, 64u, 100u, 257u, 1000u, 4099u, 20000u, 100000u};
    for (int rep = 0; rep < reps && !bad; ++rep) {
        for (uint32_t n : sizes) {
            if (n > 20000u && rep >= 3) continue;
            for (int kind = 0; kind < 8 && !bad; ++kind) {
                std::vector<float> keys(n);
                std::uniform_real_distribution<float> u(-10.f, 10.f);
                for (uint32_t i = 0; i < n; ++i) {
                    switch (kind) {
                        case 0: keys[i] = u(g); break;                                           // distinct-ish
                        case 1: keys[i] = (float) (g() % 7u); break;                             // heavy ties
                        case 2: keys[i] = (float) i; break;                                      // sorted
                        case 3: keys[i] = (float) (n - i); break;                                // reversed
                        case 4: keys[i] = 1.f; break;                                            // all equal
                        case 5: keys[i] = (float) (i < n / 2 ? i : n - i); break;                // organ pipe
                        case 6: keys[i] = (float) (g() % (n / 3u + 1u)); break;                  // pairs / triples of ties
                        default: keys[i] = (i % 3u) ? (float) i : -0.f * (float) (i & 1u); break; // zeros of both signs among sorted keys
                    }
                }
                // one segment, and the same keys cut into several segments sorted side by side
                std::vector<uint32_t> one = {0u, n};
                bad |= check(keys, one, -1, "one segment");
                std::vector<uint32_t> many = {0u};
                while (many.back() < n) many.push_back(std::min(n, many.back() + 1u + (uint32_t) (g() % (n / 3u + 2u))));
                bad |= check(keys, many, -1, "several segments");
                // depth limits that run out: heap-sorted ranges
                for (int d : {0, 1, 3}) bad |= check(keys, one, d, "depth override");
                cases += 5;
            }
        }
    }
    // partial_sort_exact alone
    for (int rep = 0; rep < 200 && !bad; ++rep) {
        const uint32_t n = 2u + (uint32_t) (g() % 300u);
        std::vector<KeyIdx> ref(n);
        std::vector<float> k(n);
        std::vector<uint32_t> idx(n);
        for (uint32_t i = 0; i < n; ++i) { k[i] = (float) (g() % (rep % 2 ? 5u : 1000u)); idx[i] = i; ref[i] = {k[i], i}; }
        auto comp = __gnu_cxx::__ops::__iter_comp_iter([](const KeyIdx &a, const KeyIdx &b) { return a.key < b.key; });
        std::__partial_sort(ref.data(), ref.data() + n, ref.data() + n, comp);
        partial_sort_exact(k.data(), idx.data(), 0, n);
        for (uint32_t i = 0; i < n; ++i)
            if (idx[i] != ref[i].idx) { std::printf("MISMATCH heap: n %u element %u\n", n, i); bad = 1; break; }
        ++cases;
    }
    std::printf("%s: %ld cases\n", bad ? "FAILED" : "ok", cases);
    return bad;
}
