// greedy_coop.cuh -- one _step() of ONE alignment by a whole warp: the end game of greedy_kernel.
//
// A pair needs 2.7 _step()s on average (150 bp, 10 % errors) but one pair in a few thousand needs 10 ... 30, each a
// chain of ~900 dependent instructions when one thread walks its 2k+1 lanes.  Inside a launch that does not matter
// (other threads carry on), but the launch cannot end before the last such pair does: a per-warp timeline of the
// 1 M-pair launch (tools/timeline_probe.py) shows 99 % of the warps gone 409 us after the start and the kernel
// ending at 472 us.  So once a warp has nothing left to hand out and only a few of its threads still hold a pair,
// the warp finishes those pairs one after the other with its threads spread over the pair's LANES: thread i
// refreshes lane i of the owner's highway table (the owner's lanes and cache sit in shared memory, any thread of the
// CTA can address them), the arg-max of _update_highway_list is three warp max-reductions over integer keys, and
// _choose_best_highway's ordered scan is a ballot loop -- the arithmetic per lane is pair_step's, line for line
// (greedy.cuh), and so are the results.  The owner keeps position, penalty and the op list.
#pragma once
#include "greedy.cuh"

namespace gasma {

#if defined(__CUDA_ARCH__) || defined(__CUDACC__)

// order-preserving unsigned key of a double (sign bit flipped, negative values complemented; -0 folded into +0)
__device__ __forceinline__ unsigned long long order_key(double h) {
    const long long bits = __double_as_longlong(dadd(h, 0.0));
    return (unsigned long long)bits ^ ((unsigned long long)(bits >> 63) | 0x8000000000000000ull);
}

// So = the OWNER's lane store; ps / sink / status are read and written in the owner thread only.  The pair's lane
// count (ub - lb + 1) must not exceed 32.  Returns true (to every thread) when the traversal is over.
template <int W, bool CORR, int FLIP>
__device__ __forceinline__ bool pair_step_coop(const KParams& P, const LaneStore<W>& So, int owner, int lane, PairState& ps,
                                               OpSink& sink, uint32_t& status) {
    constexpr int BITS = 64 * W;
    constexpr unsigned FULL = 0xffffffffu;
    using ST = typename LaneState<W>::type;
    const int x = P.x, o = P.o, e = P.e;
    const bool GLOBAL = P.global != 0;
    const int m = __shfl_sync(FULL, ps.m, owner), n = __shfl_sync(FULL, ps.n, owner);
    const int cl = __shfl_sync(FULL, ps.cl, owner), cc = __shfl_sync(FULL, ps.cc, owner);
    const int lb = CORR ? __shfl_sync(FULL, ps.lb, owner) : -P.k, ub = CORR ? __shfl_sync(FULL, ps.ub, owner) : P.k;
    const bool first = __shfl_sync(FULL, (int)ps.first, owner) != 0;
    int over = 0;
    if (lane == owner && ++ps.steps > 4 * BITS + 16) { status |= kStatusSteps; over = 1; }
    if (__shfl_sync(FULL, over, owner) != 0) return true;
    const bool leap_cost = GLOBAL || !first;
    const int dl = n - m, nl = ub - lb + 1;
    const bool on = lane < nl;
    const int l = lane + lb;

    // ---- _update_highway_list (hurdle_matrix.h:291-352): thread i refreshes lane i ----
    ST st = 0;
    int sw = 0;
    bool reach_l = false;
    unsigned long long key_s = 0ull, key_r = 0ull;   // 0 = no candidate
    uint32_t k2_s = 0u, k2_r = 0u;
    if (on) {
        const uint32_t* row = lane_base<W>(So, lane);
        const int sc = cc + sfc(cl, l);
        const int de = lane_destination(m, n, l);
        st = So.state[lane * So.stride];
        int sp = st_sp(st), len = st_len(st), nsw = st_nsw(st);
        if (sp < sc) {
            nsw = iabs(l - cl);
            scan_highway<W, FLIP>(row, So.stride, sc, de, sp, len);
            if (sp + len > de) { len = imax(0, de - sp); reach_l = true; }
        }
        sw = leap_cost ? slp(cl, l, o, e) : 0;
        const int nh = pcb<W>(row, So.stride, sc, sp + len);
        st = LaneState<W>::pack(sp, len, nsw, nh);
        So.state[lane * So.stride] = st;
        // significance heuristic (:328-332) and destination heuristic (:334-342); ties go to the larger leap, then to
        // the FIRST lane: (leap, 63 - lane) in one word.  leap = -(switch costs) lies in (-2^22, 0], 63 - lane in [13, 113]
        const double h = dadd(dadd(dmul(P.ms, (double)len), dmul(P.mms, (double)nh)), dmul(P.ids, (double)nsw));
        key_s = order_key(h);
        k2_s = ((uint32_t)((1 << 22) - sw) << 8) | (uint32_t)(63 - l);
        const int fsw = GLOBAL ? slp(l, dl, o, e) : 0;
        const int hr = -sw - x * nh - fsw - x * (de - sp - len);
        key_r = order_key((double)hr);
        k2_r = ((uint32_t)((1 << 22) - sw - fsw) << 8) | (uint32_t)(63 - l);
    }
    const bool reach = __ballot_sync(FULL, reach_l) != 0u;
    const unsigned long long hk = reach ? key_r : key_s;
    const uint32_t k2 = reach ? k2_r : k2_s;
    const uint32_t hk_hi = (uint32_t)(hk >> 32), hk_lo = (uint32_t)hk;
    const uint32_t m_hi = __reduce_max_sync(FULL, hk_hi);
    const uint32_t m_lo = __reduce_max_sync(FULL, hk_hi == m_hi ? hk_lo : 0u);
    const uint32_t m2 = __reduce_max_sync(FULL, (hk != 0ull && hk_hi == m_hi && hk_lo == m_lo) ? k2 : 0u);
    const int bl = 63 - (int)(m2 & 0xFFu);
    const ST st_bl = __shfl_sync(FULL, st, bl - lb);
    if (st_len(st_bl) <= 0) return true;   // no usable highway (:358-360, :570)

    // ---- _choose_best_highway (:368-401): lanes in ascending order, taken if not worse on both counts ----
    const int spb = st_sp(st_bl);
    int sT = x * st_nh(st_bl) + (leap_cost ? slp(cl, bl, o, e) : 0), sI = sT, b = bl;
    const uint32_t* ob = lane_base<W>(So, bl - lb);
    bool ok = on && l != bl;
    int tc = 0, ic = 0;
    if (ok) {
        const int sp2 = st_sp(st), s_l = sfc(l, bl);
        ok = sp2 + s_l <= spb;
        if (ok) {
            const int end = sp2 + st_len(st);
            ic = sw + st_nh(st);
            tc = ic + slp(l, bl, o, e) + imax(0, x * pcb<W>(ob, So.stride, s_l + end, spb));
        }
    }
    int after = -1;
    for (;;) {
        const uint32_t take = __ballot_sync(FULL, ok && lane > after && tc <= sT && ic <= sI);
        if (!take) break;
        after = __ffs((int)take) - 1;
        sT = __shfl_sync(FULL, tc, after);
        sI = __shfl_sync(FULL, ic, after);
        b = after + lb;
    }
    const ST st_b = __shfl_sync(FULL, st, b - lb);
    // ---- _step (:412-433) ----
    const int endb = st_sp(st_b) + st_len(st_b);
    if (lane == owner) {
        ps.cost += (leap_cost ? slp(cl, b, o, e) : 0) + x * st_nh(st_b);
        sink.update(b, cl, endb - (cc + sfc(cl, b)));
        ps.cl = b;
        ps.cc = endb;
        ps.first = false;
    }
    return endb >= lane_destination(m, n, b);
}

#endif

}  // namespace gasma
