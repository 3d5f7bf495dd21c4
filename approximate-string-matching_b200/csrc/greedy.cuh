// greedy.cuh -- K2 (lane construction) + K3 (greedy highway traversal, penalty, CIGAR ops).
//
// Replaces hurdle_matrix<int_128bit>::reset()+run() (GASMA/hurdle_matrix.h:625-665, :568-597) with
// its helpers _construct_hurdles (:441-455), _update_highway_list (:285-362), _choose_best_highway
// (:368-401), _step (:407-434), _update_CIGAR (:238-251), generalised to W 64-bit words per lane
// (W=2 is the reference's int_128bit).  Semantics follow SURVEY.md section 8a rows a4-a12 literally,
// including the stale per-lane cache and the popcount-as-length final op.
//
// One alignment per thread.  The 2k+1 lanes (orig + short-hurdle-flipped) and the per-lane cached
// highway live in shared memory, laid out [lane][word][thread] so that a warp's accesses are
// conflict-free; the running position, penalty and the two arg-max candidates live in registers.
#pragma once
#include "common.cuh"

namespace gasma {

// per-lane cached highway, packed: [15:0] starting_point+1, [31:16] length, [47:32] num_switches, [63:48] num_hurdles
GASMA_HD uint64_t st_pack(int sp, int len, int nsw, int nh) {
    return (uint64_t)(uint32_t)((sp + 1) | (len << 16)) | ((uint64_t)(uint32_t)(nsw | (nh << 16)) << 32);
}
GASMA_HD int st_sp(uint64_t s) { return (int)((uint32_t)s & 0xFFFFu) - 1; }
GASMA_HD int st_len(uint64_t s) { return (int)((uint32_t)s >> 16); }
GASMA_HD int st_nsw(uint64_t s) { return (int)((uint32_t)(s >> 32) & 0xFFFFu); }
GASMA_HD int st_nh(uint64_t s) { return (int)((uint32_t)(s >> 32) >> 16); }

// index of the first bit >= s that is set in (v ^ inv); -1 if none   (first_one/first_zero, utils.h:168-191)
template <int W>
GASMA_HD int scan_from(const uint64_t v[W], uint64_t inv, int s) {
    int res = -1;
#pragma unroll
    for (int j = W - 1; j >= 0; --j) {
        uint64_t t = (v[j] ^ inv) & ones_from(s, j);
        if (t) res = 64 * j + ctz64(t);
    }
    return res;
}

// pop_count_between(from,to), utils.h:263-270 = shift_left(from).shift_right(from+BITS-to).pop_count():
// bits [from, to) clipped to the vector, 0 if from >= to, 0 if the second shift would be negative.
template <int W>
GASMA_HD int pcb(const uint64_t v[W], int from, int to) {
    if (to > 64 * W + from) return 0;
    int c = 0;
#pragma unroll
    for (int j = 0; j < W; ++j) c += popc64(v[j] & ones_from(from, j) & ~ones_from(to, j));
    return c;
}

// v >> s (towards bit 0), zero fill, for a small constant-ish shift 0 <= s < 64   (shift_left, utils.h:143)
template <int W>
GASMA_HD void shift_down_small(const uint64_t v[W], int s, uint64_t out[W]) {
#pragma unroll
    for (int j = 0; j < W; ++j) {
        uint64_t hi = (j + 1 < W) ? v[j + 1] : 0ull;
        out[j] = s ? (shr64(v[j], s) | shl64(hi, 64 - s)) : v[j];
    }
}

// One lane of _construct_hurdles (hurdle_matrix.h:441-455) + flip_short_hurdles(1) (utils.h:200-216)
template <int W>
GASMA_HD void build_lane(const uint64_t A0[W], const uint64_t A1[W], const uint64_t B0[W], const uint64_t B1[W],
                         int lane, uint64_t orig[W], uint64_t flip[W]) {
    uint64_t s0[W], s1[W];
    const int a = iabs(lane);
    if (lane < 0) {
        shift_down_small<W>(A0, a, s0);
        shift_down_small<W>(A1, a, s1);
#pragma unroll
        for (int j = 0; j < W; ++j) orig[j] = (s0[j] ^ B0[j]) | (s1[j] ^ B1[j]);
    } else {
        shift_down_small<W>(B0, a, s0);
        shift_down_small<W>(B1, a, s1);
#pragma unroll
        for (int j = 0; j < W; ++j) orig[j] = (s0[j] ^ A0[j]) | (s1[j] ^ A1[j]);
    }
#pragma unroll
    for (int j = 0; j < W; ++j) {
        uint64_t up = (orig[j] << 1) | (j ? orig[j - 1] >> 63 : 0ull);
        uint64_t dn = (orig[j] >> 1) | (j + 1 < W ? orig[j + 1] << 63 : 0ull);
        flip[j] = orig[j] & (up | dn);
    }
}

// Strided per-thread view of the lane storage: element (lane index li, kind, word w) of this thread.
// kind 0 = orig, 1 = flip.  `stride` is the number of threads sharing the array (1 in the host harness).
struct LaneStore {
    uint64_t* lanes;  // [(li*2+kind)*W + w][stride]
    uint64_t* state;  // [li][stride]
    int stride;
};

template <int W>
GASMA_HD void lane_load(const LaneStore& S, int li, int kind, uint64_t v[W]) {
#pragma unroll
    for (int j = 0; j < W; ++j) v[j] = S.lanes[((li * 2 + kind) * W + j) * S.stride];
}
template <int W>
GASMA_HD void lane_store(const LaneStore& S, int li, int kind, const uint64_t v[W]) {
#pragma unroll
    for (int j = 0; j < W; ++j) S.lanes[((li * 2 + kind) * W + j) * S.stride] = v[j];
}

struct OpSink {
    uint16_t* ops;   // this pair's staging row (may be null: costs only)
    int cap;
    int n;
    GASMA_HD void emit(int code, int len) {
        if (ops && n < cap) ops[n] = (uint16_t)((len << 3) | code);
        ++n;
    }
    // _update_CIGAR, hurdle_matrix.h:238-251
    GASMA_HD void update(int best, int cur, int total) {
        if (best < cur) emit(1, cur - best);
        else if (best > cur) emit(2, best - cur);
        if (total > 0) emit(0, total);
    }
};

// The whole alignment of one pair.  Returns status bits (0 = fine).
template <int W>
GASMA_HD uint32_t greedy_pair(const KParams& P, const uint64_t A0[W], const uint64_t A1[W], const uint64_t B0[W],
                              const uint64_t B1[W], int m, int n, const LaneStore& S, OpSink& sink, int& cost_out,
                              int& steps_out) {
    constexpr int BITS = 64 * W;
    const int k = P.k, x = P.x, o = P.o, e = P.e;
    const bool GLOBAL = P.global != 0;

    // reset(): highways::reset (:106-119) + _construct_hurdles (:441-455)
    for (int l = -k; l <= k; ++l) {
        uint64_t og[W], fl[W];
        build_lane<W>(A0, A1, B0, B1, l, og, fl);
        lane_store<W>(S, l + k, 0, og);
        lane_store<W>(S, l + k, 1, fl);
        S.state[(l + k) * S.stride] = st_pack(-1, 0, BITS, BITS);
    }
    const int dl = n - m;
    int cl = 0, cc = 0, cost = 0, steps = 0;
    bool first = true;
    uint32_t status = 0;

    for (;;) {
        if (++steps > 4 * BITS + 16) { status |= kStatusSteps; break; }
        // ---- _update_highway_list, loops 1 and 2 fused: both arg-max candidates are tracked, the
        // one that applies is picked once `reach` is known (hurdle_matrix.h:291-352) ----
        bool reach = false;
        double bestH_s = 0.0; int bestLeap_s = 0, bl_s = 0; uint64_t st_s = 0; bool have_s = false;
        int bestH_r = 0, bestLeap_r = 0, bl_r = 0; uint64_t st_r = 0; bool have_r = false;
        for (int l = -k; l <= k; ++l) {
            const int li = l + k;
            const int sc = cc + sfc(cl, l);
            const int de = lane_destination(m, n, l);
            uint64_t st = S.state[li * S.stride];
            int sp = st_sp(st), len = st_len(st), nsw = st_nsw(st);
            if (sp < sc) {  // cache miss: look up the closest highway of this lane (:293-310)
                nsw = iabs(l - cl);
                uint64_t f[W];
                lane_load<W>(S, li, 1, f);
                int p0 = scan_from<W>(f, ~0ull, sc);             // first zero at/after sc
                int pos0 = p0 >= 0 ? p0 : imax(sc, BITS);        // zero fill beyond the vector
                int p1 = scan_from<W>(f, 0ull, pos0);            // next hurdle
                int nx = p1 >= 0 ? p1 - pos0 : BITS;             // first_one of an empty vector = BITS
                sp = pos0;
                len = nx;
                if (sp + len > de) { len = imax(0, de - sp); reach = true; }
            }
            const int sw = (GLOBAL || !first) ? slp(cl, l, o, e) : 0;
            uint64_t og[W];
            lane_load<W>(S, li, 0, og);
            const int nh = pcb<W>(og, sc, sp + len);
            const int hc = x * nh;
            st = st_pack(sp, len, nsw, nh);
            S.state[li * S.stride] = st;
            // significance heuristic (:328-332): ((ms*len) + (mms*nh)) + (ids*nsw), each op rounded
            const double h = dadd(dadd(dmul(P.ms, (double)len), dmul(P.mms, (double)nh)), dmul(P.ids, (double)nsw));
            const int lh = -sw;
            if (!have_s || h > bestH_s || (h == bestH_s && lh > bestLeap_s)) {
                have_s = true; bestH_s = h; bestLeap_s = lh; bl_s = l; st_s = st;
            }
            // destination heuristic (:334-342), integer valued
            const int fsw = GLOBAL ? slp(l, dl, o, e) : 0;
            const int hr = -sw - hc - fsw - x * (de - sp - len);
            const int lhr = -sw - fsw;
            if (!have_r || hr > bestH_r || (hr == bestH_r && lhr > bestLeap_r)) {
                have_r = true; bestH_r = hr; bestLeap_r = lhr; bl_r = l; st_r = st;
            }
        }
        const int bl = reach ? bl_r : bl_s;
        const uint64_t st_bl = reach ? st_r : st_s;
        if (st_len(st_bl) <= 0) break;  // no usable highway (:358-360)

        // ---- _choose_best_highway (:368-401) ----
        const int spb = st_sp(st_bl);
        const int swb = (GLOBAL || !first) ? slp(cl, bl, o, e) : 0;
        int sT = x * st_nh(st_bl) + swb, sI = sT, b = bl;
        uint64_t st_b = st_bl;
        uint64_t ob[W];
        lane_load<W>(S, bl + k, 0, ob);
        for (int l = -k; l <= k; ++l) {
            if (l == bl) continue;
            const uint64_t st = S.state[(l + k) * S.stride];
            const int sp = st_sp(st), s_l = sfc(l, bl);
            if (sp + s_l > spb) continue;
            const int end = sp + st_len(st);
            const int swl = (GLOBAL || !first) ? slp(cl, l, o, e) : 0;
            const int ic = swl + st_nh(st);  // == switch_cost + pop_count_between(cc+sfc(cl,l), end), no *x (:388)
            const int hp = x * pcb<W>(ob, s_l + end, spb);
            const int tc = ic + slp(l, bl, o, e) + imax(0, hp);
            if (tc <= sT && ic <= sI) { sT = tc; sI = ic; b = l; st_b = st; }
        }
        // ---- _step (:412-433) ----
        const int swc = (GLOBAL || !first) ? slp(cl, b, o, e) : 0;
        cost += swc + x * st_nh(st_b);
        const int endb = st_sp(st_b) + st_len(st_b);
        sink.update(b, cl, endb - (cc + sfc(cl, b)));
        cl = b;
        cc = endb;
        first = false;
        if (cc >= lane_destination(m, n, cl)) break;
    }
    // ---- run() fix-up (:575-590).  A destination lane outside the band keeps the state of a fresh
    // object: lane value 1 (utils.h:57-59) and destination -|lane| (hurdle_matrix.h:79-92 with m=n=0) ----
    {
        const bool in_band = iabs(dl) <= k;
        const int dc = in_band ? lane_destination(m, n, dl) : -iabs(dl);
        if (cl != dl || cc < dc) {
            uint64_t od[W];
            if (in_band) lane_load<W>(S, dl + k, 0, od);
            else {
#pragma unroll
                for (int j = 0; j < W; ++j) od[j] = j ? 0ull : 1ull;
            }
            const int d = pcb<W>(od, cc + sfc(cl, dl), dc);
            cost += (GLOBAL ? slp(cl, dl, o, e) : 0) + imax(0, x * d);
            sink.update(dl, cl, d);
        }
    }
    cost_out = cost;
    steps_out = steps;
    return status;
}

}  // namespace gasma
