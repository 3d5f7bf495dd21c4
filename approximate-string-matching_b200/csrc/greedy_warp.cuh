// greedy_warp.cuh -- K2+K3 for long reads and wide bands: ONE ALIGNMENT PER WARP.
//
// Same algorithm and same per-lane arithmetic as greedy.cuh (hurdle_matrix.h:285-597 generalised to W words), but
// the 2k+1 lanes of one alignment are spread over the 32 threads of a warp (lane index li -> thread li % 32), the
// way north_star (3) describes: every thread refreshes the highways of its own lanes, `reaching_destination` is a
// __ballot_sync, the longest highway is picked by a __shfl_xor_sync arg-max over (heuristic, leap, lane), and
// _choose_best_highway's sequential "take if not worse" scan becomes a ballot loop that jumps from taken lane to
// taken lane.  Needed where one thread cannot hold an alignment: 1 kbp reads are 16 words per lane and k = 50 is
// 101 lanes (12.9 KB of hurdle masks per alignment, kept in the warp's slice of shared memory).
#pragma once
#include "greedy.cuh"
#include "kernels.cuh"

namespace gasma {

template <int W>
struct WarpLayout {
    static constexpr int NW = 2 * W;              // 32-bit words per lane
    static constexpr int PITCH = NW + 1;          // odd pitch: threads reading the same word of 32 lanes hit 32 banks
    static constexpr int PLANE = NW + 2;          // plane words + zero padding for shifted reads
    static constexpr int STREAM = ((4 * W + 2) + 2 * 2 + 7) / 8 * 8;   // uint16 chunks of one string (+ window slack)
    __host__ __device__ static constexpr size_t bytes_per_warp(int nl) {   // nl = lane slots per alignment (KParams::nl)
        return (size_t)nl * PITCH * 4 + (size_t)nl * 8 + 4 * PLANE * 4 + 4 * STREAM * 2 + 16;
    }
};

// candidate of the arg-max: key (heuristic, leap), lane, cached highway state
struct WarpCand {
    double h;       // significance heuristic, or the integer destination heuristic converted exactly
    int leap, l;
    uint64_t st;
    bool valid;
};
// the reference keeps the FIRST lane (ascending) among equal (heuristic, leap) keys: strict '>' while scanning upwards
__device__ __forceinline__ bool cand_better(const WarpCand& a, const WarpCand& b) {
    if (a.valid != b.valid) return a.valid;
    if (!a.valid) return false;
    if (a.h != b.h) return a.h > b.h;
    if (a.leap != b.leap) return a.leap > b.leap;
    return a.l < b.l;
}
__device__ __forceinline__ WarpCand cand_shfl_xor(const WarpCand& c, int d) {
    WarpCand r;
    r.h = __shfl_xor_sync(0xffffffffu, c.h, d);
    r.leap = __shfl_xor_sync(0xffffffffu, c.leap, d);
    r.l = __shfl_xor_sync(0xffffffffu, c.l, d);
    r.st = __shfl_xor_sync(0xffffffffu, c.st, d);
    r.valid = __shfl_xor_sync(0xffffffffu, (int)c.valid, d) != 0;
    return r;
}

// 32 bits of a uint16-chunk stream starting at bit `bit`, for plane word w of a string of `len` characters
__device__ __forceinline__ uint32_t stream_word(const uint32_t* stream32, int bit, int len, int w) {
    const int pos = bit + 32 * w, idx = pos >> 5, sh = pos & 31;
    const uint32_t v = __funnelshift_r(stream32[idx], stream32[idx + 1], sh);
    const int keep = len - 32 * w;                       // bits of this word inside the string ('A' tail beyond)
    return keep >= 32 ? v : (keep <= 0 ? 0u : (v & (0xFFFFFFFFu >> (32 - keep))));
}

template <int W, int WARPS>
__global__ void __launch_bounds__(32 * WARPS)
greedy_warp_kernel(KParams P, const uint8_t* __restrict__ reads, const uint32_t* __restrict__ read_off,
                   const uint8_t* __restrict__ refs, const uint32_t* __restrict__ ref_off, long long n,
                   int32_t* __restrict__ cost, uint16_t* __restrict__ ops_rows, uint16_t* __restrict__ ops_ovf, int ops_stride,
                   uint32_t* __restrict__ nops_out, uint32_t* __restrict__ pair_counter,
                   uint32_t* __restrict__ status, unsigned long long* __restrict__ step_counter,
                   uint32_t* __restrict__ tile_sums) {
    using L = WarpLayout<W>;
    constexpr int BITS = 64 * W, NW = L::NW, PITCH = L::PITCH;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int k = P.k, nl_cap = P.nl, x = P.x, o = P.o, e = P.e;
    const bool GLOBAL = P.global != 0;
    unsigned char* mine = smem_raw + (size_t)warp * ((L::bytes_per_warp(nl_cap) + 15) / 16 * 16);
    uint64_t* s_state = reinterpret_cast<uint64_t*>(mine);
    uint32_t* s_lanes = reinterpret_cast<uint32_t*>(s_state + nl_cap);
    uint32_t* s_plane = s_lanes + (size_t)nl_cap * PITCH;        // [4][PLANE]: A0, A1, B0, B1
    uint16_t* s_stream = reinterpret_cast<uint16_t*>(s_plane + 4 * L::PLANE);   // [4][STREAM]
    for (int i = lane; i < 4 * L::STREAM; i += 32) s_stream[i] = 0;
    const uint32_t rtotal = read_off[n], ftotal = ref_off[n];
    uint32_t st_bits = 0;
    unsigned long long steps_total = 0;

    for (;;) {
        uint32_t pi = 0;
        if (lane == 0) pi = atomicAdd(pair_counter, 1u);
        pi = __shfl_sync(0xffffffffu, pi, 0);
        if ((long long)pi >= n) break;
        const long long p = pi;
        const uint32_t r0 = read_off[p], r1 = read_off[p + 1], f0 = ref_off[p], f1 = ref_off[p + 1];
        const int m = (int)(r1 - r0), nn = (int)(f1 - f0);
        uint32_t bad = 0;
        if (m > BITS || nn > BITS) bad |= kStatusLength;
        else if (iabs(nn - m) > kMaxK) bad |= kStatusBand;
        else if (P.correction && (iabs(nn - m) + k > kMaxK || 2 * k + 1 + iabs(nn - m) > nl_cap)) bad |= kStatusBand;
        if (bad) {
            st_bits |= bad;
            if (lane == 0) { cost[p] = INT32_MIN; if (nops_out) nops_out[p] = 0u; }
            continue;
        }
        // ---- K1: pack both strings (coalesced 16-byte chunks -> 2-bit streams -> plane words) ----
        __syncwarp();
        const uint32_t rbase = r0 & ~15u, fbase = f0 & ~15u;
        stage_tile(reads, rbase, r1, rtotal, s_stream, s_stream + L::STREAM, 4 * W + 2, lane, 32);
        stage_tile(refs, fbase, f1, ftotal, s_stream + 2 * L::STREAM, s_stream + 3 * L::STREAM, 4 * W + 2, lane, 32);
        __syncwarp();
        for (int i = lane; i < 4 * L::PLANE; i += 32) {
            const int pl = i / L::PLANE, w = i - pl * L::PLANE;
            uint32_t v = 0u;
            if (w < NW)
                v = stream_word(reinterpret_cast<const uint32_t*>(s_stream + pl * L::STREAM), (int)((pl < 2 ? r0 - rbase : f0 - fbase)),
                                pl < 2 ? m : nn, w);
            s_plane[i] = v;
        }
        __syncwarp();
        // ---- K2: lanes (_construct_hurdles, hurdle_matrix.h:441-455) and highway cache reset (:106-119) ----
        // lane range of this pair: -k..k, or the widened band of the -DCORRECTION build (hurdle_matrix.h:634-641)
        int lb = -k, ub = k;
        if (P.correction) { if (m <= nn) ub = nn - m + k; else lb = nn - m - k; }
        const int nl = ub - lb + 1;
        for (int li = lane; li < nl; li += 32) {
            const int l = li + lb, a = iabs(l), aw = a >> 5, ab = a & 31;
            const uint32_t* X0 = s_plane + (l < 0 ? 0 : 2) * L::PLANE;   // the shifted string: read for l<0, ref for l>=0
            const uint32_t* X1 = X0 + L::PLANE;
            const uint32_t* Y0 = s_plane + (l < 0 ? 2 : 0) * L::PLANE;
            const uint32_t* Y1 = Y0 + L::PLANE;
            uint32_t* row = s_lanes + (size_t)li * PITCH;
            // |l| <= MAX_K = 50, so aw <= 1 and the last word read is PLANE-1 (the planes carry two zero words on top);
            // each plane word is loaded once and slides from "upper" to "lower" half of the funnel shift
            uint32_t x0a = X0[aw], x1a = X1[aw];
            for (int w = 0; w < NW; ++w) {
                const uint32_t x0b = X0[w + aw + 1], x1b = X1[w + aw + 1];
                row[w] = (__funnelshift_r(x0a, x0b, ab) ^ Y0[w]) | (__funnelshift_r(x1a, x1b, ab) ^ Y1[w]);
                x0a = x0b; x1a = x1b;
            }
            s_state[li] = st_pack(-1, 0, BITS, BITS);
        }
        __syncwarp();

        // ---- K3: traversal ----
        const int dl = nn - m;
        int cl = 0, cc = 0, pen = 0, steps = 0, n_ops = 0;
        bool first = true;
        // lane 0 keeps the pair's op list (primary row / overflow row, greedy.cuh:OpSink)
        uint16_t* my_ops = nops_out ? ops_rows + (size_t)p * kRowPitch : nullptr;   // nops_out == nullptr: costs only
        uint16_t* my_ovf = nops_out ? ops_ovf + (size_t)p * ops_stride : nullptr;
        auto emit = [&](int code, int len) {
            if (lane == 0 && my_ops && n_ops < ops_stride) (n_ops < kRowOps ? my_ops : my_ovf)[n_ops] = (uint16_t)((len << 3) | code);
            ++n_ops;
        };
        auto update_cigar = [&](int best, int cur, int total) {   // _update_CIGAR, hurdle_matrix.h:238-251
            if (best < cur) emit(1, cur - best);
            else if (best > cur) emit(2, best - cur);
            if (total > 0) emit(0, total);
        };
        for (;;) {
            if (++steps > 4 * BITS + 16) { st_bits |= kStatusSteps; break; }
            const bool leap_cost = GLOBAL || !first;
            // _update_highway_list (:291-352): every thread refreshes its own lanes and keeps its best candidates
            bool reach_local = false;
            WarpCand cs{0.0, 0, 0, 0ull, false}, cr{0.0, 0, 0, 0ull, false};
            for (int li = lane; li < nl; li += 32) {
                const int l = li + lb;
                const uint32_t* row = s_lanes + (size_t)li * PITCH;
                const int sc = cc + sfc(cl, l);
                const int de = lane_destination(m, nn, l);
                uint64_t stv = s_state[li];
                int sp = st_sp(stv), len = st_len(stv), nsw = st_nsw(stv);
                if (sp < sc) {
                    nsw = iabs(l - cl);
                    scan_highway<W>(row, 1, sc, de, sp, len);
                    if (sp + len > de) { len = imax(0, de - sp); reach_local = true; }
                }
                const int sw = leap_cost ? slp(cl, l, o, e) : 0;
                const int nh = pcb<W>(row, 1, sc, sp + len);
                stv = st_pack(sp, len, nsw, nh);
                s_state[li] = stv;
                const double h = dadd(dadd(dmul(P.ms, (double)len), dmul(P.mms, (double)nh)), dmul(P.ids, (double)nsw));
                WarpCand c{h, -sw, l, stv, true};
                if (cand_better(c, cs)) cs = c;
                const int fsw = GLOBAL ? slp(l, dl, o, e) : 0;
                WarpCand d{(double)(-sw - x * nh - fsw - x * (de - sp - len)), -sw - fsw, l, stv, true};
                if (cand_better(d, cr)) cr = d;
            }
            const bool reach = __ballot_sync(0xffffffffu, reach_local) != 0u;
            const WarpCand mine = reach ? cr : cs;
            // Warp arg-max of (heuristic, leap, -lane) in three integer max-reductions (REDUX) instead of five
            // rounds of 7-register shuffles: the double is mapped to an unsigned key of the same order (sign bit
            // flipped, negative values complemented; -0 folded into +0 first), its high and low words are reduced
            // one after the other, then (leap, -lane) among the lanes that tie.  leap = -(switch costs) lies in
            // (-2^22, 0] for penalties <= 1000, 63 - lane in [13, 113].
            unsigned long long hk = 0ull;
            if (mine.valid) {
                const long long bits = __double_as_longlong(dadd(mine.h, 0.0));
                hk = (unsigned long long)bits ^ ((unsigned long long)(bits >> 63) | 0x8000000000000000ull);
            }
            const uint32_t hk_hi = (uint32_t)(hk >> 32), hk_lo = (uint32_t)hk;
            const uint32_t m_hi = __reduce_max_sync(0xffffffffu, hk_hi);
            const uint32_t m_lo = __reduce_max_sync(0xffffffffu, hk_hi == m_hi ? hk_lo : 0u);
            const bool tie = mine.valid && hk_hi == m_hi && hk_lo == m_lo;
            const uint32_t k2 = tie ? (((uint32_t)(mine.leap + (1 << 22)) << 8) | (uint32_t)(63 - mine.l)) : 0u;
            const uint32_t m2 = __reduce_max_sync(0xffffffffu, k2);
            __syncwarp();   // every lane's refreshed state is visible before it is read back / _choose_best_highway reads it
            WarpCand best;
            best.l = 63 - (int)(m2 & 0xFFu);
            best.st = s_state[best.l - lb];
            const int bl = best.l;
            if (st_len(best.st) <= 0) break;   // no usable highway (:358-360)

            // _choose_best_highway (:368-401)
            const int spb = st_sp(best.st);
            int sT = x * st_nh(best.st) + (leap_cost ? slp(cl, bl, o, e) : 0), sI = sT, b = bl;
            uint64_t st_b = best.st;
            const uint32_t* ob = s_lanes + (size_t)(bl - lb) * PITCH;
            for (int base = 0; base < nl; base += 32) {
                const int li = base + lane, l = li + lb;
                bool ok = li < nl && l != bl;
                int tc = 0, ic = 0;
                uint64_t stv = 0;
                if (ok) {
                    stv = s_state[li];
                    const int sp = st_sp(stv), s_l = sfc(l, bl);
                    ok = sp + s_l <= spb;
                    if (ok) {
                        const int end = sp + st_len(stv);
                        ic = (leap_cost ? slp(cl, l, o, e) : 0) + st_nh(stv);
                        tc = ic + slp(l, bl, o, e) + imax(0, x * pcb<W>(ob, 1, s_l + end, spb));
                    }
                }
                int after = -1;
                for (;;) {
                    const uint32_t take = __ballot_sync(0xffffffffu, ok && lane > after && tc <= sT && ic <= sI);
                    if (!take) break;
                    after = __ffs((int)take) - 1;
                    sT = __shfl_sync(0xffffffffu, tc, after);
                    sI = __shfl_sync(0xffffffffu, ic, after);
                    b = base + after + lb;
                }
            }
            if (b != bl) st_b = s_state[b - lb];   // the taken lane's highway (its state was refreshed above)
            // _step (:412-433)
            pen += (leap_cost ? slp(cl, b, o, e) : 0) + x * st_nh(st_b);
            const int endb = st_sp(st_b) + st_len(st_b);
            update_cigar(b, cl, endb - (cc + sfc(cl, b)));
            cl = b;
            cc = endb;
            first = false;
            if (cc >= lane_destination(m, nn, cl)) break;
        }
        // run() fix-up (:575-590)
        {
            const bool in_band = dl >= lb && dl <= ub;
            const int dc = in_band ? lane_destination(m, nn, dl) : -iabs(dl);
            if (cl != dl || cc < dc) {
                int d = 0;
                if (in_band) d = pcb<W>(s_lanes + (size_t)(dl - lb) * PITCH, 1, cc + sfc(cl, dl), dc);
                pen += (GLOBAL ? slp(cl, dl, o, e) : 0) + imax(0, x * d);
                update_cigar(dl, cl, d);
            }
        }
        if (lane == 0) {
            if (my_ops && n_ops > kRowKeep)   // consumers read lists of more than kRowKeep ops from the overflow row
                for (int i = 0; i < (n_ops < kRowOps ? n_ops : kRowOps); ++i) my_ovf[i] = my_ops[i];
            cost[p] = pen;
            if (nops_out) nops_out[p] = (uint32_t)n_ops;
            if (tile_sums && n_ops) atomicAdd(tile_sums + (p >> kCsrTileShift), (uint32_t)n_ops);
            if (n_ops > ops_stride) st_bits |= kStatusSteps;
            steps_total += (unsigned long long)steps;
        }
        __syncwarp();
    }
    const uint32_t wst = __reduce_or_sync(0xffffffffu, st_bits);
    if (wst && lane == 0) atomicOr(status, wst);
    if (step_counter && lane == 0 && steps_total) atomicAdd(step_counter, steps_total);
}

}  // namespace gasma
