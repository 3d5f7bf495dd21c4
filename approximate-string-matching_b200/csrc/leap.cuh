// leap.cuh -- the benchmark's LEAP comparator on the device: banded affine Landau-Vishkin.
//
// Replaces class LV (GASMA/benchmark/LEAP_SIMD/LV_BAG.cpp) as benchmark::_run_LEAP drives it
// (GASMA/benchmark/benchmark_utils.h:156-178) on an object set up with init(k, 200, ED_GLOBAL, x, o, e) (:289):
// the value the benchmark records is get_ED() = the error level at which the best converging lane reaches the
// end of the longer string (LV_BAG.cpp:236-246,360) -- only that penalty feeds the report's LEAP accuracy line.
//
// The reference extends a lane character by character (count_ID_length, :9-23).  Here a lane is the same kind of
// mismatch bit mask the greedy aligner builds (one string shifted against the other, XOR of the 2-bit planes), so
// an extension is a find-first-set from the start column.  Lane d = l - (k+1): d < 0 offsets the read by |d|,
// d > 0 the reference; both strings are NUL-padded to L = max(m, n) in the reference (load_reads, :112-123), i.e.
// every column at or past lim_d = min(m + max(0,-d), n + max(0,d)) mismatches -- those mask bits are set.
// Error levels only look back max(x, o, e) levels, so the tables are rings of that depth (+1) instead of
// [lane][201]; a cell that is not reached is written as -2 each level, which is what an untouched cell of the
// reference holds.  Needs o >= e (the reference's own assumption, :165) and max(x, o) <= kLeapMaxDepth - 1.
#pragma once
#include "common.cuh"

namespace gasma {

constexpr int kLeapAf = 200;     // af_threshold of the benchmark's LV object (benchmark_utils.h:289)
constexpr int kLeapMaxDepth = 8; // deepest ring: error levels q-7 .. q

// Shared-memory view of one alignment: word i of any table lives at base[i * stride].
struct LeapStore {
    uint32_t* lanes;   // [(2k+1) * NW]  mismatch masks, lane d at index (d + k) * NW
    int* en;           // [(2k+3) * depth]  end of the matching run
    int* ip;           // [(2k+3) * depth]  best start reached through an insertion
    int* dp;           // [(2k+3) * depth]  ... through a deletion
    int stride;
    int depth;         // ring depth: a power of two > max(x, o, e)
};
GASMA_HD int leap_depth(int x, int o, int e) {
    const int need = imax(x, imax(o, e)) + 1;
    int d = 2;
    while (d < need) d <<= 1;
    return d;
}
GASMA_HD int leap_words_per_pair(int k, int nw, int depth) { return (2 * k + 1) * nw + 3 * (2 * k + 3) * depth; }

// first mismatch column at or after pos on one lane (count_ID_length): pos >= L returns pos
template <int NW>
GASMA_HD int leap_extend(const uint32_t* lane, int stride, int pos, int L) {
    if (pos >= L) return pos;
    int w = pos >> 5;
    uint32_t v = lane[w * stride] & (0xFFFFFFFFu << (pos & 31));
    while (v == 0u) {
        if (++w >= NW) return L;
        v = lane[w * stride];
    }
#if defined(__CUDA_ARCH__)
    const int p = 32 * w + __ffs((int)v) - 1;
#else
    const int p = 32 * w + __builtin_ctz(v);
#endif
    return p < L ? p : L;
}

// v << s over W words (towards higher columns), 0 <= s < 64
template <int W>
GASMA_HD void leap_shift_up(const uint64_t v[W], int s, uint64_t out[W]) {
#pragma unroll
    for (int j = 0; j < W; ++j) {
        const uint64_t lo = j > 0 ? v[j - 1] : 0ull;
        out[j] = s ? (shl64(v[j], s) | shr64(lo, 64 - s)) : v[j];
    }
}

// Returns the LEAP penalty, or INT32_MIN when no lane converges within kLeapAf error levels.
template <int W>
GASMA_HD int leap_pair(int k, int x, int o, int e, const uint64_t A0[W], const uint64_t A1[W], const uint64_t B0[W],
                       const uint64_t B1[W], int m, int n, const LeapStore& S) {
    constexpr int NW = 2 * W;
    const int total = 2 * k + 3, mid = k + 1, L = m > n ? m : n, st = S.stride, kLeapDepth = S.depth;
    // lane masks
    for (int d = -k; d <= k; ++d) {
        uint64_t s0[W], s1[W];
        const int a = d < 0 ? -d : d;
        const int lim = d < 0 ? imin(m + a, n) : imin(m, n + a);
        if (d < 0) { leap_shift_up<W>(A0, a, s0); leap_shift_up<W>(A1, a, s1); }
        else { leap_shift_up<W>(B0, a, s0); leap_shift_up<W>(B1, a, s1); }
        uint32_t* dst = S.lanes + (size_t)((d + k) * NW) * st;
#pragma unroll
        for (int j = 0; j < W; ++j) {
            uint64_t mm = d < 0 ? ((s0[j] ^ B0[j]) | (s1[j] ^ B1[j])) : ((s0[j] ^ A0[j]) | (s1[j] ^ A1[j]));
            const int left = lim - 64 * j;                       // columns of this word still inside both strings
            if (left <= 0) mm = ~0ull;
            else if (left < 64) mm |= ~0ull << left;
            dst[(2 * j) * st] = (uint32_t)mm;
            dst[(2 * j + 1) * st] = (uint32_t)(mm >> 32);
        }
    }
    // level 0: only the main lane is live (ED_GLOBAL, LV_BAG.cpp:99-101)
    for (int l = 0; l < total; ++l) {
        S.en[(l * kLeapDepth) * st] = -2;
        S.ip[(l * kLeapDepth) * st] = -2;
        S.dp[(l * kLeapDepth) * st] = -2;
    }
    {
        const int end0 = leap_extend<NW>(S.lanes + (size_t)(k * NW) * st, st, 0, L);
        if (end0 == L) return 0;
        S.en[(mid * kLeapDepth) * st] = end0;
    }
    for (int q = 1; q <= kLeapAf; ++q) {
        const int cur = q & (kLeapDepth - 1);
        const int so = (q - o) & (kLeapDepth - 1), se = (q - e) & (kLeapDepth - 1), sx = (q - x) & (kLeapDepth - 1);
        bool pass = false;
        // the sentinel lanes are never reached
        S.en[(0 * kLeapDepth + cur) * st] = -2; S.ip[(0 * kLeapDepth + cur) * st] = -2; S.dp[(0 * kLeapDepth + cur) * st] = -2;
        S.en[((total - 1) * kLeapDepth + cur) * st] = -2; S.ip[((total - 1) * kLeapDepth + cur) * st] = -2;
        S.dp[((total - 1) * kLeapDepth + cur) * st] = -2;
        for (int l = 1; l < total - 1; ++l) {
            const int top = l >= mid, bot = l <= mid;
            // insertion: from the lane below (LV_BAG.cpp:166-178)
            int iv = -2;
            {
                const int eo = q >= o ? S.en[((l - 1) * kLeapDepth + so) * st] : -2;
                const int ie = q >= e ? S.ip[((l - 1) * kLeapDepth + se) * st] : -2;
                if (q >= o && eo >= 0 && eo > ie) iv = eo + top;
                else if (ie >= 0) iv = ie + top;
            }
            // deletion: from the lane above (:180-183)
            int dv = -2;
            {
                const int eo = q >= o ? S.en[((l + 1) * kLeapDepth + so) * st] : -2;
                const int de = q >= e ? S.dp[((l + 1) * kLeapDepth + se) * st] : -2;
                if (q >= o && eo >= 0 && eo > de) dv = eo + bot;
                else if (de >= 0) dv = de + bot;
            }
            int s0 = -2;
            if (q >= x) {
                const int ex = S.en[(l * kLeapDepth + sx) * st];
                if (ex >= 0) s0 = ex + 1;
            }
            if (iv > s0) s0 = iv;
            if (dv > s0) s0 = dv;
            int end = -2;
            if (s0 >= 0) {
                end = leap_extend<NW>(S.lanes + (size_t)((l - 1) * NW) * st, st, s0, L);
                if (end == L) {
                    const int d = iabs(mid - l);
                    if (q + (d ? o + (d - 1) * e : 0) <= kLeapAf) pass = true;   // :236-246: the level is what is reported
                }
            }
            S.ip[(l * kLeapDepth + cur) * st] = iv;
            S.dp[(l * kLeapDepth + cur) * st] = dv;
            S.en[(l * kLeapDepth + cur) * st] = end;
        }
        if (pass) return q;
    }
    return INT32_MIN;
}

}  // namespace gasma
