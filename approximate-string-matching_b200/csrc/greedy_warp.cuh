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
    // a lane row = one zero word, NW mask words, three zero words: the highway search reads word w-1 ... w+3 of its
    // position without bounds tests.  Odd pitch: threads reading the same word of 32 lanes hit 32 banks
    static constexpr int PAD_LO = 1, PAD_HI = 3;
    static constexpr int PITCH = (NW + PAD_LO + PAD_HI) | 1;
    static constexpr int PLANE = NW + 2;          // plane words + zero padding for shifted reads
    static constexpr int STREAM = ((4 * W + 2) + 2 * 2 + 7) / 8 * 8;   // uint16 chunks of one string (+ window slack)
    __host__ __device__ static constexpr size_t bytes_per_warp(int nl) {   // nl = lane slots per alignment (KParams::nl)
        return (size_t)nl * PITCH * 4 + (size_t)nl * 8 + 4 * PLANE * 4 + 4 * STREAM * 2 + 16;
    }
};

// The closest highway at/after column sc (greedy.cuh:scan_highway, same results) on a zero-padded row: `row` points at
// mask word 0, row[-1] and row[NW .. NW+2] are zero.  The shipped mask variant only (flip_short_hurdles(1)).
template <int W>
__device__ __forceinline__ void scan_highway_padded(const uint32_t* row, int sc, int de, int& sp, int& len) {
    constexpr int NW = 2 * W, BITS = 64 * W;
    const int w = sc >> 5, b = sc & 31;
    if (w >= NW) { sp = imax(sc, BITS); len = BITS; return; }
    const uint32_t om = row[w - 1], o0 = row[w], o1 = row[w + 1], o2 = row[w + 2], o3 = row[w + 3];
    const uint32_t f0 = o0 & (funnel_l1(om, o0) | funnel_r1(o0, o1)), f1 = o1 & (funnel_l1(o0, o1) | funnel_r1(o1, o2)),
                   f2 = o2 & (funnel_l1(o1, o2) | funnel_r1(o2, o3));
    const uint32_t lo = __funnelshift_r(f0, f1, b), hi = __funnelshift_r(f1, f2, b);   // flipped lane bits [sc, sc+64)
    const uint32_t zlo = ~lo, zhi = ~hi;
    if ((zlo | zhi) == 0u) { scan_highway_loop<W, 0>(row, 1, sc, sp, len); return; }   // 64 hurdles in a row
    const int fz = zlo ? ctz32(zlo) : 32 + ctz32(zhi);
    const uint32_t tlo = fz < 32 ? (lo & (0xFFFFFFFFu << fz)) : 0u;
    const uint32_t thi = fz < 32 ? hi : (hi & (0xFFFFFFFFu << (fz - 32)));
    sp = sc + fz;
    if ((tlo | thi) != 0u) len = (tlo ? ctz32(tlo) : 32 + ctz32(thi)) - fz;
    else len = (sc + 64 >= BITS || sc + 64 > de) ? BITS : scan_highway_tail<W, 0>(row, 1, sc + 64, de, sp);
}

// 32 bits of a uint16-chunk stream starting at bit `bit`, for plane word w of a string of `len` characters
__device__ __forceinline__ uint32_t stream_word(const uint32_t* stream32, int bit, int len, int w) {
    const int pos = bit + 32 * w, idx = pos >> 5, sh = pos & 31;
    const uint32_t v = __funnelshift_r(stream32[idx], stream32[idx + 1], sh);
    const int keep = len - 32 * w;                       // bits of this word inside the string ('A' tail beyond)
    return keep >= 32 ? v : (keep <= 0 ? 0u : (v & (0xFFFFFFFFu >> (32 - keep))));
}


// ITER = lane slots per thread (ceil(lane slots / 32)): the per-slot loops are unrolled and every thread keeps the
// highway cache of its own lanes (and this step's switch cost) in registers between _update_highway_list and
// _choose_best_highway; a copy in shared memory is what the warp reads the winner's cache from.
template <int W, int WARPS, int ITER>
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
    uint64_t* s_state = reinterpret_cast<uint64_t*>(mine);       // [nl_cap]: every lane's cached highway, for the warp to read
    uint32_t* s_lanes = reinterpret_cast<uint32_t*>(s_state + nl_cap);   // [nl_cap][PITCH], mask word 0 of a row at PAD_LO
    uint32_t* s_plane = s_lanes + (size_t)nl_cap * PITCH;        // [4][PLANE]: A0, A1, B0, B1
    uint16_t* s_stream = reinterpret_cast<uint16_t*>(s_plane + 4 * L::PLANE);   // [4][STREAM]
    for (int i = lane; i < 4 * L::STREAM; i += 32) s_stream[i] = 0;
    for (int i = lane; i < nl_cap * PITCH; i += 32) s_lanes[i] = 0u;   // the rows' zero padding, written once
    const uint32_t* my_row0 = s_lanes + (size_t)lane * PITCH + L::PAD_LO;   // mask word 0 of this thread's first lane
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
        const int l0 = lane + lb;                  // this thread's lanes: l0, l0 + 32, ...
        uint64_t stv[ITER];                        // their cached highways
        int sw_[ITER];                             // their switch cost in the current step
#pragma unroll
        for (int i = 0; i < ITER; ++i) {
            stv[i] = st_pack(-1, 0, BITS, BITS);
            sw_[i] = 0;
        }
        for (int li = lane; li < nl; li += 32) {
            const int l = li + lb, a = iabs(l), aw = a >> 5, ab = a & 31;
            const uint32_t* X0 = s_plane + (l < 0 ? 0 : 2) * L::PLANE;   // the shifted string: read for l<0, ref for l>=0
            const uint32_t* X1 = X0 + L::PLANE;
            const uint32_t* Y0 = s_plane + (l < 0 ? 2 : 0) * L::PLANE;
            const uint32_t* Y1 = Y0 + L::PLANE;
            uint32_t* row = s_lanes + (size_t)li * PITCH + L::PAD_LO;
            // |l| <= MAX_K = 50, so aw <= 1 and the last word read is PLANE-1 (the planes carry two zero words on top);
            // each plane word is loaded once and slides from "upper" to "lower" half of the funnel shift
            uint32_t x0a = X0[aw], x1a = X1[aw];
            for (int w = 0; w < NW; ++w) {
                const uint32_t x0b = X0[w + aw + 1], x1b = X1[w + aw + 1];
                row[w] = (__funnelshift_r(x0a, x0b, ab) ^ Y0[w]) | (__funnelshift_r(x1a, x1b, ab) ^ Y1[w]);
                x0a = x0b; x1a = x1b;
            }
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
            // _update_highway_list (:291-352): every thread refreshes its own lanes and keeps its best candidate as an
            // integer key: hk = the FP64 significance heuristic mapped to an unsigned of the same order, k2 = (leap, -lane).
            // The reference keeps the FIRST lane (ascending) among equal (heuristic, leap): 63 - lane in the low bits.
            // leap = -(switch costs) lies in (-2^22, 0] for penalties <= 1000, 63 - lane in [13, 113].
            bool reach_local = false;
            unsigned long long hk = 0ull;   // 0 = no candidate
            uint32_t k2 = 0u;
#pragma unroll
            for (int i = 0; i < ITER; ++i) {
                if (lane + 32 * i < nl) {
                    const int l = l0 + 32 * i;
                    const uint32_t* row = my_row0 + (size_t)(32 * i) * PITCH;
                    const int sc = cc + sfc(cl, l);
                    int sp = st_sp(stv[i]), len = st_len(stv[i]), nsw = st_nsw(stv[i]);
                    if (sp < sc) {
                        const int de = lane_destination(m, nn, l);
                        nsw = iabs(l - cl);
                        scan_highway_padded<W>(row, sc, de, sp, len);
                        if (sp + len > de) { len = imax(0, de - sp); reach_local = true; }
                    }
                    const int sw = leap_cost ? slp(cl, l, o, e) : 0;
                    const int nh = pcb<W>(row, 1, sc, sp + len);
                    sw_[i] = sw;
                    stv[i] = st_pack(sp, len, nsw, nh);
                    s_state[lane + 32 * i] = stv[i];
                    const double h = dadd(dadd(dmul(P.ms, (double)len), dmul(P.mms, (double)nh)), dmul(P.ids, (double)nsw));
                    const unsigned long long hki = order_key(h);
                    const uint32_t k2i = ((uint32_t)((1 << 22) - sw) << 8) | (uint32_t)(63 - l);
                    if (hki > hk || (hki == hk && k2i > k2)) { hk = hki; k2 = k2i; }
                }
            }
            if (__ballot_sync(0xffffffffu, reach_local) != 0u) {
                // a highway reaches its destination: the candidates are ranked by the destination heuristic instead
                // (:327-347; rare -- normally the last step of an alignment)
                hk = 0ull;
                k2 = 0u;
#pragma unroll
                for (int i = 0; i < ITER; ++i) {
                    if (lane + 32 * i < nl) {
                        const int l = l0 + 32 * i, de = lane_destination(m, nn, l);
                        const int fsw = GLOBAL ? slp(l, dl, o, e) : 0;
                        const int hv = -sw_[i] - x * st_nh(stv[i]) - fsw - x * (de - st_sp(stv[i]) - st_len(stv[i]));
                        const unsigned long long hki = order_key((double)hv);
                        const uint32_t k2i = ((uint32_t)((1 << 22) - sw_[i] - fsw) << 8) | (uint32_t)(63 - l);
                        if (hki > hk || (hki == hk && k2i > k2)) { hk = hki; k2 = k2i; }
                    }
                }
            }
            // warp arg-max in three integer max-reductions (REDUX): high word, low word, then (leap, -lane) among the ties
            const uint32_t hk_hi = (uint32_t)(hk >> 32), hk_lo = (uint32_t)hk;
            const uint32_t m_hi = __reduce_max_sync(0xffffffffu, hk_hi);
            const uint32_t m_lo = __reduce_max_sync(0xffffffffu, hk_hi == m_hi ? hk_lo : 0u);
            const uint32_t m2 = __reduce_max_sync(0xffffffffu, (hk != 0ull && hk_hi == m_hi && hk_lo == m_lo) ? k2 : 0u);
            const int bl = 63 - (int)(m2 & 0xFFu);
            __syncwarp();   // every lane's refreshed cache is visible
            const uint64_t st_best = s_state[bl - lb];
            if (st_len(st_best) <= 0) break;   // no usable highway (:358-360)

            // _choose_best_highway (:368-401): lanes are visited in ascending order and taken if not worse on both
            // counts; both bounds only ever go down, so a lane that fails them before its hurdle count is known is out
            const int spb = st_sp(st_best);
            int sT = x * st_nh(st_best) + (leap_cost ? slp(cl, bl, o, e) : 0), sI = sT, b = bl;
            const uint32_t* ob = s_lanes + (size_t)(bl - lb) * PITCH + L::PAD_LO;
#pragma unroll
            for (int i = 0; i < ITER; ++i) {
                if (32 * i < nl) {
                    const int l = l0 + 32 * i;
                    bool ok = lane + 32 * i < nl && l != bl;
                    int tc = 0, ic = 0;
                    if (ok) {
                        const int sp = st_sp(stv[i]), s_l = sfc(l, bl);
                        ic = sw_[i] + st_nh(stv[i]);
                        const int lo = ic + slp(l, bl, o, e);
                        ok = sp + s_l <= spb && ic <= sI && lo <= sT;
                        if (ok) tc = lo + imax(0, x * pcb<W>(ob, 1, s_l + sp + st_len(stv[i]), spb));
                    }
                    int after = -1;
                    for (;;) {
                        const uint32_t take = __ballot_sync(0xffffffffu, ok && lane > after && tc <= sT && ic <= sI);
                        if (!take) break;
                        after = __ffs((int)take) - 1;
                        sT = __shfl_sync(0xffffffffu, tc, after);
                        sI = __shfl_sync(0xffffffffu, ic, after);
                        b = 32 * i + after + lb;
                    }
                }
            }
            const uint64_t st_b = b != bl ? s_state[b - lb] : st_best;   // the taken lane's highway
            __syncwarp();   // read by everyone before the next step's owners overwrite it
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
                if (in_band) d = pcb<W>(s_lanes + (size_t)(dl - lb) * PITCH + L::PAD_LO, 1, cc + sfc(cl, dl), dc);
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
