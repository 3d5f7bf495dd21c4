// oracle/gasma_fast.cpp -- TEST / MEASUREMENT INFRASTRUCTURE, NOT PRODUCT CODE.
//
// The CPU arm behind bench.py's `cpu_baseline` / `--impl reference` for workloads the unmodified reference cannot
// run (strings longer than its 128-character vectors): a CPU implementation of the greedy path written for speed,
// the way the reference's own int_128bit code is -- W fixed at compile time, bit-vectors as W 64-bit words in
// registers, no per-operation structs or memsets, BMI/POPCNT (-march=x86-64-v3), FP64 without contraction.
// The checker port (gasma_oracle.c) is written for clarity and is several times slower; a speed-up quoted against it
// would be inflated.  Pinned: bit-equal to gasma_oracle.c at W = 1..4 and to the unmodified reference (oracle/_ref)
// at W = 2 (tests/test_oracle.py::test_fast_cpu_arm_*); at W = 2 it must not be slower than the unmodified reference.
// Only tests/ and bench.py's CPU legs load it.
//
// Semantics: SURVEY.md section 8a rows a1-a12 / Appendix A (hurdle_matrix.h:285-597, utils.h:49-271, 576-593),
// canonical mode ('A' tails).  All file:line citations are relative to /root/reference.
#include <cmath>
#include <cstdint>
#include <cstring>
#include <thread>
#include <vector>
#include <x86intrin.h>

namespace {

constexpr int kMaxK = 50;   // hurdle_matrix.h:8

struct Params {
    int k, x, o, e, global;
    double ms, mms, ids;   // hurdle_matrix.h:536-538
};

template <int W>
struct Vec {
    uint64_t w[W];
};

// bit_convert.cpp:248-369 for one string, 8 characters per step: ASCII A=0x41 C=0x43 G=0x47 T=0x54 -> bit 2 is the
// G/T plane, bit 1 ^ bit 2 the C/T plane; a block that holds any other byte takes the exact per-character route
// (such bytes pack as 00).
template <int W>
inline void pack(const uint8_t* s, int len, Vec<W>& p0, Vec<W>& p1) {
    for (int j = 0; j < W; ++j) p0.w[j] = p1.w[j] = 0;
    int i = 0;
    for (; i + 8 <= len; i += 8) {
        uint64_t x;
        memcpy(&x, s + i, 8);
        const uint64_t h = x & 0x0202020202020202ull, g = x & 0x0404040404040404ull;
        const uint64_t t = (g >> 2) & ~(h >> 1);
        const uint64_t rebuilt = 0x4141414141414141ull + (h | g) + 15ull * t;
        uint64_t b1, b2;
        if (rebuilt == x) {
            b1 = _pext_u64(x, 0x0202020202020202ull);
            b2 = _pext_u64(x, 0x0404040404040404ull);
            b1 ^= b2;
        } else {
            b1 = b2 = 0;
            for (int q = 0; q < 8; ++q) {
                const uint8_t c = s[i + q];
                b1 |= (uint64_t)(c == 'C' || c == 'T') << q;
                b2 |= (uint64_t)(c == 'G' || c == 'T') << q;
            }
        }
        p0.w[i >> 6] |= b1 << (i & 63);
        p1.w[i >> 6] |= b2 << (i & 63);
    }
    for (; i < len; ++i) {
        const uint8_t c = s[i];
        p0.w[i >> 6] |= (uint64_t)(c == 'C' || c == 'T') << (i & 63);
        p1.w[i >> 6] |= (uint64_t)(c == 'G' || c == 'T') << (i & 63);
    }
}

// logical shift towards bit 0 by 0 <= s < 64 (the reference's "shift_left", utils.h:143-153)
template <int W>
inline Vec<W> shr(const Vec<W>& v, int s) {
    Vec<W> r;
    if (s == 0) return v;
    for (int j = 0; j < W; ++j) r.w[j] = (v.w[j] >> s) | (j + 1 < W ? v.w[j + 1] << (64 - s) : 0);
    return r;
}
// flip_short_hurdles(1), utils.h:200-216: a hurdle survives iff a neighbouring bit is a hurdle too
template <int W>
inline Vec<W> flip1(const Vec<W>& v) {
    Vec<W> r;
    for (int j = 0; j < W; ++j) {
        const uint64_t up = (v.w[j] >> 1) | (j + 1 < W ? v.w[j + 1] << 63 : 0);
        const uint64_t dn = (v.w[j] << 1) | (j > 0 ? v.w[j - 1] >> 63 : 0);
        r.w[j] = v.w[j] & (up | dn);
    }
    return r;
}
// first bit at/after `from` whose value is `one` (after xor with inv); 64W if none
template <int W>
inline int next_bit(const Vec<W>& v, int from, uint64_t inv) {
    int j = from >> 6;
    if (j >= W) return 64 * W;
    uint64_t t = (v.w[j] ^ inv) & (~0ull << (from & 63));
    for (;;) {
        if (t) return 64 * j + (int)_tzcnt_u64(t);
        if (++j >= W) return 64 * W;
        t = v.w[j] ^ inv;
    }
}
// pop_count_between(from, to), utils.h:263-270: bits [from, to) clipped to the vector; 0 if from >= to or if the
// second shift (from + BITS - to) would be negative
template <int W>
inline int pcb(const Vec<W>& v, int from, int to) {
    constexpr int BITS = 64 * W;
    if (to > BITS + from) return 0;
    if (to > BITS) to = BITS;
    if (from >= to) return 0;
    const int j0 = from >> 6, j1 = (to - 1) >> 6;
    const uint64_t mlo = ~0ull << (from & 63), mhi = ~0ull >> (63 - ((to - 1) & 63));
    if (j0 == j1) return (int)_mm_popcnt_u64(v.w[j0] & mlo & mhi);
    int c = (int)_mm_popcnt_u64(v.w[j0] & mlo) + (int)_mm_popcnt_u64(v.w[j1] & mhi);
    for (int j = j0 + 1; j < j1; ++j) c += (int)_mm_popcnt_u64(v.w[j]);
    return c;
}

inline int iabs(int a) { return a < 0 ? -a : a; }
inline int imax(int a, int b) { return a > b ? a : b; }
inline int imin(int a, int b) { return a < b ? a : b; }
// utils.h:576-579, :587-593; hurdle_matrix.h:58-68
inline int slp(int l1, int l2, int o, int e) { return l1 == l2 ? 0 : o + e * (iabs(l1 - l2) - 1); }
inline int sfc(int l1, int l2) { return ((long long)l1 * l2 >= 0) ? imax(0, iabs(l1) - iabs(l2)) : iabs(l1); }
inline int dest(int m, int n, int l) { return imin(m + imin(l, 0), n - imax(l, 0)); }

struct Ops {
    std::vector<uint16_t>* out;
    int n;
    inline void emit(int code, int len) {
        if (out) out->push_back((uint16_t)((len << 3) | code));
        ++n;
    }
    // _update_CIGAR, hurdle_matrix.h:238-251
    inline void update(int best, int cur, int total) {
        if (best < cur) emit(1, cur - best);
        else if (best > cur) emit(2, best - cur);
        if (total > 0) emit(0, total);
    }
};

// reset() + run() + get_cost() + get_CIGAR() of one pair (hurdle_matrix.h:625-665, :568-597), Appendix A of SURVEY.md.
// Returns the penalty, INT32_MIN if the pair is outside the domain (length > 64W, |n-m| > MAX_K).
template <int W>
int align(const uint8_t* read, int m, const uint8_t* ref, int n, const Params& P, Ops& ops, long long& steps_total) {
    constexpr int BITS = 64 * W;
    const int k = P.k, x = P.x, o = P.o, e = P.e;
    const bool GLOBAL = P.global != 0;
    if (m > BITS || n > BITS || iabs(n - m) > kMaxK) return INT32_MIN;
    Vec<W> A0, A1, B0, B1;
    pack<W>(read, m, A0, A1);
    pack<W>(ref, n, B0, B1);
    const int nl = 2 * k + 1;
    Vec<W> orig[2 * kMaxK + 1], flip[2 * kMaxK + 1];
    int sp[2 * kMaxK + 1], len[2 * kMaxK + 1], nsw[2 * kMaxK + 1], nh[2 * kMaxK + 1], sw[2 * kMaxK + 1], de[2 * kMaxK + 1];
    for (int li = 0; li < nl; ++li) {   // _construct_hurdles :441-455, highways::reset :106-119
        const int l = li - k, a = iabs(l);
        Vec<W> v;
        if (l < 0) {
            const Vec<W> s0 = shr<W>(A0, a), s1 = shr<W>(A1, a);
            for (int j = 0; j < W; ++j) v.w[j] = (s0.w[j] ^ B0.w[j]) | (s1.w[j] ^ B1.w[j]);
        } else {
            const Vec<W> s0 = shr<W>(B0, a), s1 = shr<W>(B1, a);
            for (int j = 0; j < W; ++j) v.w[j] = (s0.w[j] ^ A0.w[j]) | (s1.w[j] ^ A1.w[j]);
        }
        orig[li] = v;
        flip[li] = flip1<W>(v);
        sp[li] = -1; len[li] = 0; nsw[li] = nh[li] = sw[li] = BITS;
        de[li] = dest(m, n, l);
    }
    const int dl = n - m;
    int cl = 0, cc = 0, cost = 0, steps = 0;
    bool first = true;
    for (;;) {
        if (++steps > 4 * BITS + 16) break;   // cannot happen; the port and the kernels carry the same guard
        const bool leap_cost = GLOBAL || !first;
        bool reach = false;
        for (int li = 0; li < nl; ++li) {     // _update_highway_list loop 1, :291-326
            const int l = li - k;
            const int sc = cc + sfc(cl, l);
            if (sp[li] < sc) {
                nsw[li] = iabs(l - cl);
                const int s = imax(sc, next_bit<W>(flip[li], sc, ~0ull));     // first non-hurdle at/after sc
                const int q = next_bit<W>(flip[li], s, 0ull);                 // next hurdle; first_one(0) = BITS
                sp[li] = s;
                len[li] = q >= BITS ? BITS : q - s;
                if (s + len[li] > de[li]) { len[li] = imax(0, de[li] - s); reach = true; }
            }
            sw[li] = leap_cost ? slp(cl, l, o, e) : 0;
            nh[li] = pcb<W>(orig[li], sc, sp[li] + len[li]);
        }
        double bestH = -INFINITY;
        int bestLeap = 0, bl = 0;              // best_leap starts at 0 (:287)
        for (int li = 0; li < nl; ++li) {     // loop 2, :328-352
            const int l = li - k;
            double h = ((P.ms * (double)len[li]) + (P.mms * (double)nh[li])) + (P.ids * (double)nsw[li]);
            int lh = -sw[li];
            if (reach) {
                const int f = GLOBAL ? slp(l, dl, o, e) : 0;
                h = (double)(-sw[li] - x * nh[li] - f - x * (de[li] - sp[li] - len[li]));
                lh -= f;
            }
            if (h > bestH || (h == bestH && lh > bestLeap)) { bestH = h; bestLeap = lh; bl = l; }
        }
        if (len[bl + k] <= 0) break;           // no usable highway (:358-360)
        const int spb = sp[bl + k];            // _choose_best_highway, :368-401
        int sT = x * nh[bl + k] + sw[bl + k], sI = sT, b = bl;
        for (int li = 0; li < nl; ++li) {
            const int l = li - k;
            if (l == bl) continue;
            const int s_l = sfc(l, bl);
            if (sp[li] + s_l > spb) continue;
            const int end = sp[li] + len[li];
            const int ic = sw[li] + nh[li];    // switch_cost + pop_count_between(cc + sfc(cl,l), end): no *x (:388)
            const int tc = ic + slp(l, bl, o, e) + imax(0, x * pcb<W>(orig[bl + k], s_l + end, spb));
            if (tc <= sT && ic <= sI) { sT = tc; sI = ic; b = l; }
        }
        cost += sw[b + k] + x * nh[b + k];     // _step, :412-433
        const int endb = sp[b + k] + len[b + k];
        ops.update(b, cl, endb - (cc + sfc(cl, b)));
        cl = b;
        cc = endb;
        first = false;
        if (cc >= de[cl + k]) break;
    }
    {   // run() fix-up, :575-590; a destination lane outside the band is a fresh object's lane: value 1, destination -|lane|
        const bool in_band = dl >= -k && dl <= k;
        const int dc = in_band ? de[dl + k] : -iabs(dl);
        if (cl != dl || cc < dc) {
            const int d = in_band ? pcb<W>(orig[dl + k], cc + sfc(cl, dl), dc) : 0;
            cost += (GLOBAL ? slp(cl, dl, o, e) : 0) + imax(0, x * d);
            ops.update(dl, cl, d);
        }
    }
    steps_total += steps;
    return cost;
}

struct Shard {
    std::vector<uint16_t> ops;
    long long steps = 0;
};

template <int W>
void run_range(const uint8_t* reads, const uint32_t* roff, const uint8_t* refs, const uint32_t* foff, int64_t lo, int64_t hi,
               const Params& P, int32_t* cost, uint32_t* nops, Shard* sh) {
    if (nops) sh->ops.reserve((size_t)(hi - lo) * 8);
    for (int64_t i = lo; i < hi; ++i) {
        Ops ops{nops ? &sh->ops : nullptr, 0};
        cost[i] = align<W>(reads + roff[i], (int)(roff[i + 1] - roff[i]), refs + foff[i], (int)(foff[i + 1] - foff[i]), P, ops, sh->steps);
        if (nops) nops[i] = (uint32_t)ops.n;
    }
}

}  // namespace

extern "C" {

// Batch over CSR strings, `threads` contiguous shards.  cost[n]; op_off[n+1] and ops (CSR, same encoding as
// include/gasma_b200.h) are optional -- op_off without ops still builds every CIGAR (the work the reference does) and
// reports the offsets.  Returns 0, -1 bad argument, -4 ops_cap too small (op_off is complete then).
int gf_greedy_batch(const uint8_t* reads, const uint32_t* roff, const uint8_t* refs, const uint32_t* foff, int64_t n,
                    int k, int x, int o, int e, int global, int W, double pm, double pmm, double pid, int threads,
                    int32_t* cost, uint16_t* ops, uint32_t* op_off, int64_t ops_cap, int64_t* total_steps) {
    if (n < 0 || !cost || k < 0 || k > kMaxK || W < 1 || W > 4) return -1;
    const Params P{k, x, o, e, global, std::log(pm / 0.25), std::log(pmm / 0.25), std::log(pid / 2 / 0.25)};
    if (threads < 1) threads = 1;
    if (threads > n) threads = n > 0 ? (int)n : 1;
    std::vector<Shard> sh(threads);
    std::vector<uint32_t> nops(op_off ? (size_t)n : 0);
    auto work = [&](int t) {
        const int64_t lo = n * t / threads, hi = n * (t + 1) / threads;
        uint32_t* np = op_off ? nops.data() : nullptr;
        switch (W) {
            case 1: run_range<1>(reads, roff, refs, foff, lo, hi, P, cost, np, &sh[t]); break;
            case 2: run_range<2>(reads, roff, refs, foff, lo, hi, P, cost, np, &sh[t]); break;
            case 3: run_range<3>(reads, roff, refs, foff, lo, hi, P, cost, np, &sh[t]); break;
            default: run_range<4>(reads, roff, refs, foff, lo, hi, P, cost, np, &sh[t]); break;
        }
    };
    if (threads == 1) work(0);
    else {
        std::vector<std::thread> pool;
        for (int t = 0; t < threads; ++t) pool.emplace_back(work, t);
        for (auto& th : pool) th.join();
    }
    long long steps = 0;
    for (auto& s : sh) steps += s.steps;
    if (total_steps) *total_steps = steps;
    if (!op_off) return 0;
    uint64_t acc = 0;
    for (int64_t i = 0; i < n; ++i) { op_off[i] = (uint32_t)acc; acc += nops[i]; }
    op_off[n] = (uint32_t)acc;
    if (!ops) return 0;
    if ((int64_t)acc > ops_cap) return -4;
    for (int t = 0; t < threads; ++t) {
        const int64_t lo = n * t / threads;
        if (!sh[t].ops.empty()) memcpy(ops + op_off[lo], sh[t].ops.data(), sh[t].ops.size() * 2);
    }
    return 0;
}

}  // extern "C"
