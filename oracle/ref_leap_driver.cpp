// oracle/ref_leap_driver.cpp -- TEST INFRASTRUCTURE (checker), never linked into the product.
//
// Thin extern "C" driver around the UNMODIFIED LEAP comparator of the reference's benchmark
// (GASMA/benchmark/LEAP_SIMD/LV_BAG.{h,cpp}: class LV, banded affine Landau-Vishkin), compiled where the
// sources lie under /root/reference (see oracle/Makefile; output goes to oracle/_ref/).  Nothing is copied.
//
// The call sequence is benchmark::_run_LEAP (GASMA/benchmark/benchmark_utils.h:156-178) with the object set up
// as the benchmark does (:286,:289): init(k, 200, ED_GLOBAL, x, o, e); per pair
//   load_reads(s1, s2, max(len1, len2)); reset(); run(); if (check_pass()) backtrack(); get_ED();
// mode 0: one LV object per thread reused for all its pairs (what the benchmark does);
// mode 1: a fresh object per pair (shows that results do not depend on earlier pairs).
#include <algorithm>
#include <cstdint>
#include <cstring>
#include <thread>
#include <vector>

#include "LV_BAG.h"

namespace {

void leap_range(const char* reads, const uint32_t* read_off, const char* refs, const uint32_t* ref_off, int64_t lo, int64_t hi,
                int k, int x, int o, int e, int mode, int32_t* penalty, uint8_t* pass) {
    LV* obj = nullptr;
    if (mode == 0) { obj = new LV; obj->init(k, 200, ED_GLOBAL, x, o, e); }
    std::vector<char> s1(_MAX_LENGTH_ + 1), s2(_MAX_LENGTH_ + 1);
    for (int64_t i = lo; i < hi; ++i) {
        const int m = (int)(read_off[i + 1] - read_off[i]), n = (int)(ref_off[i + 1] - ref_off[i]);
        if (m > _MAX_LENGTH_ || n > _MAX_LENGTH_) { penalty[i] = INT32_MIN; if (pass) pass[i] = 0; continue; }
        memcpy(s1.data(), reads + read_off[i], m); s1[m] = '\0';   // the benchmark hands over std::string::c_str()
        memcpy(s2.data(), refs + ref_off[i], n); s2[n] = '\0';
        if (mode == 1) { obj = new LV; obj->init(k, 200, ED_GLOBAL, x, o, e); }
        obj->load_reads(s1.data(), s2.data(), std::max(m, n));
        obj->reset();
        obj->run();
        const bool ok = obj->check_pass();
        if (ok) obj->backtrack();
        penalty[i] = ok ? obj->get_ED() : INT32_MIN;   // a failed pair would report the previous pair's value
        if (pass) pass[i] = ok ? 1 : 0;
        if (mode == 1) delete obj;
    }
    if (mode == 0) delete obj;
}

}  // namespace

extern "C" int gasma_ref_leap_batch(const char* reads, const uint32_t* read_off, const char* refs, const uint32_t* ref_off,
                                    int64_t n, int k, int x, int o, int e, int mode, int threads, int32_t* penalty,
                                    uint8_t* pass) {
    if (threads <= 1) {
        leap_range(reads, read_off, refs, ref_off, 0, n, k, x, o, e, mode, penalty, pass);
        return 0;
    }
    std::vector<std::thread> pool;
    for (int t = 0; t < threads; ++t)
        pool.emplace_back(leap_range, reads, read_off, refs, ref_off, n * t / threads, n * (t + 1) / threads, k, x, o, e, mode,
                          penalty, pass);
    for (auto& th : pool) th.join();
    return 0;
}
