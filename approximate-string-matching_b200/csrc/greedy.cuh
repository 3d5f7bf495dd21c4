// greedy.cuh -- K2 (lane construction) + K3 (greedy highway traversal, penalty, CIGAR ops).
//
// Replaces hurdle_matrix<int_128bit>::reset()+run() (GASMA/hurdle_matrix.h:625-665, :568-597) with
// its helpers _construct_hurdles (:441-455), _update_highway_list (:285-362), _choose_best_highway
// (:368-401), _step (:407-434), _update_CIGAR (:238-251), generalised to W 64-bit words per lane
// (W=2 is the reference's int_128bit).  Semantics follow SURVEY.md section 8a rows a4-a12 literally,
// including the stale per-lane cache and the popcount-as-length final op.
//
// One alignment per thread.  The 2k+1 lanes (original hurdle masks as 32-bit words) and the per-lane cached
// highway live in shared memory, laid out [lane][word][thread] so that a warp's accesses are
// conflict-free; the running position, penalty and the two arg-max candidates live in registers.
// The traversal is split into pair_init / pair_step / pair_finish so that a kernel can interleave pairs:
// a thread that finishes takes a new pair while its warp neighbours keep stepping (kernels.cuh).
#pragma once
#include "common.cuh"

namespace gasma {

// per-lane cached highway, packed: [15:0] starting_point+1, [31:16] length, [47:32] num_switches, [63:48] num_hurdles
GASMA_HD uint64_t st_pack(int sp, int len, int nsw, int nh) {
    return (uint64_t)(uint32_t)((sp + 1) | (len << 16)) | ((uint64_t)(uint32_t)(nsw | (nh << 16)) << 32);
}
// The same four fields in ONE 32-bit word, a byte each, for vectors of up to 192 bits (W <= 3): starting_point+1 <=
// 64W+MAX_K+1 = 243, length and num_hurdles <= 192, num_switches <= 100.  Halves the state's shared memory (more
// alignments in flight per SM) and the registers / selects of the two arg-max candidates.
GASMA_HD uint32_t st_pack32(int sp, int len, int nsw, int nh) {
    return (uint32_t)((sp + 1) | (len << 8) | (nsw << 16) | (nh << 24));
}
GASMA_HD int st_sp(uint32_t s) { return (int)(s & 0xFFu) - 1; }
GASMA_HD int st_len(uint32_t s) { return (int)((s >> 8) & 0xFFu); }
GASMA_HD int st_nsw(uint32_t s) { return (int)((s >> 16) & 0xFFu); }
GASMA_HD int st_nh(uint32_t s) { return (int)(s >> 24); }
template <int W> struct LaneState { using type = uint64_t; static GASMA_HD type pack(int sp, int len, int nsw, int nh) { return st_pack(sp, len, nsw, nh); } };
template <> struct LaneState<1> { using type = uint32_t; static GASMA_HD type pack(int sp, int len, int nsw, int nh) { return st_pack32(sp, len, nsw, nh); } };
template <> struct LaneState<2> : LaneState<1> {};
template <> struct LaneState<3> : LaneState<1> {};
GASMA_HD int st_sp(uint64_t s) { return (int)((uint32_t)s & 0xFFFFu) - 1; }
GASMA_HD int st_len(uint64_t s) { return (int)((uint32_t)s >> 16); }
GASMA_HD int st_nsw(uint64_t s) { return (int)((uint32_t)(s >> 32) & 0xFFFFu); }
GASMA_HD int st_nh(uint64_t s) { return (int)((uint32_t)(s >> 32) >> 16); }

// ---------------------------------------------------------------------------------------------
// Lane storage.  A lane is kept as NW = 2W 32-bit words so that a scan or a ranged popcount touches only
// the one to three words it needs (a dynamic word index is free in shared memory, it is not in registers).
// Word w of lane index li of this thread lives at lanes[(li*NW + w) * stride]: consecutive threads are
// consecutive 4-byte words, so a warp never bank-conflicts whatever w each thread picks.  Only the ORIGINAL
// hurdle masks are stored; the short-hurdle-flipped view (flip_short_hurdles(1), utils.h:200-216) is derived
// word by word inside the scans (3 ALU ops per word), which halves the shared memory per alignment.
// `stride` is the number of threads sharing the array (1 in the host harness).
// ---------------------------------------------------------------------------------------------
template <int W>
struct LaneStore {
    uint32_t* lanes;
    typename LaneState<W>::type* state;  // [li][stride]
    int stride;
};

GASMA_HD int ctz32(uint32_t x) {  // x != 0
#if defined(__CUDA_ARCH__)
    return __ffs((int)x) - 1;
#else
    return __builtin_ctz(x);
#endif
}
GASMA_HD int popc32(uint32_t x) {
#if defined(__CUDA_ARCH__)
    return __popc(x);
#else
    return __builtin_popcount(x);
#endif
}

GASMA_HD uint32_t funnel_l1(uint32_t lo, uint32_t hi) {  // bits of (hi:lo) << 1, upper word
#if defined(__CUDA_ARCH__)
    return __funnelshift_l(lo, hi, 1);
#else
    return (hi << 1) | (lo >> 31);
#endif
}
GASMA_HD uint32_t funnel_r1(uint32_t lo, uint32_t hi) {  // bits of (hi:lo) >> 1, lower word
#if defined(__CUDA_ARCH__)
    return __funnelshift_r(lo, hi, 1);
#else
    return (lo >> 1) | (hi << 31);
#endif
}
GASMA_HD uint32_t funnel_l2(uint32_t lo, uint32_t hi) {  // bits of (hi:lo) << 2, upper word
#if defined(__CUDA_ARCH__)
    return __funnelshift_l(lo, hi, 2);
#else
    return (hi << 2) | (lo >> 30);
#endif
}
GASMA_HD uint32_t funnel_r2(uint32_t lo, uint32_t hi) {  // bits of (hi:lo) >> 2, lower word
#if defined(__CUDA_ARCH__)
    return __funnelshift_r(lo, hi, 2);
#else
    return (lo >> 2) | (hi << 30);
#endif
}
// One word of the lane the highway search runs on (hurdle_matrix.h:453), from three neighbouring words of the original
// mask.  FLIP selects the aligner's mask variant at compile time (utils.h:200-240):
//   0  flip_short_hurdles(1): a hurdle survives iff a neighbouring bit is a hurdle too (edges see 0) -- as shipped
//   1  flip_short_hurdles(2): ... iff a bit at distance 1 or 2 is a hurdle
//   2  flip_short_hurdles(1).flip_short_matches(1): then a non-hurdle between two hurdles becomes a hurdle; at the
//      ends of the vector the missing neighbour counts as a hurdle (shift_left_one / shift_right_one feed a 1,
//      utils.h:155-163).  The neighbour words' edge bits of the intermediate mask are rebuilt from prev / next;
//      first / last say whether `cur` is word 0 / the top word of the vector (only read by variant 2).
template <int FLIP>
GASMA_HD uint32_t flip_word(uint32_t prev, uint32_t cur, uint32_t next, bool first, bool last) {
    if (FLIP == 1) return cur & (funnel_l1(prev, cur) | funnel_r1(cur, next) | funnel_l2(prev, cur) | funnel_r2(cur, next));
    const uint32_t f = cur & (funnel_l1(prev, cur) | funnel_r1(cur, next));
    if (FLIP != 2) return f;
    // bit 31 of the intermediate mask's previous word, bit 0 of its next word (1 beyond the ends of the vector)
    const uint32_t fp = first ? 1u : ((prev >> 31) & ((prev >> 30) | cur) & 1u);
    const uint32_t fn = last ? 1u : (next & ((cur >> 31) | (next >> 1)) & 1u);
    const uint32_t below = (f << 1) | fp;          // bit i = intermediate bit i-1
    const uint32_t above = (f >> 1) | (fn << 31);  // bit i = intermediate bit i+1
    return f | (below & above);
}

// The closest highway of a lane at/after column sc (hurdle_matrix.h:296-303 on the flipped lane):
//   sp  = sc + first_zero(flip >> sc)          first non-hurdle at/after sc; max(sc, BITS) if there is none
//   len = first_one(flip >> sp)                run of non-hurdles from sp; BITS if no hurdle follows (utils.h:168-182)
// One loop serves both searches (it flips from "looking for a 0" to "looking for a 1" in place and carries on in
// the same word), touching only the words between sc and the end of the highway.
// `base` points at word 0 of the lane's original mask for this thread.
template <int W, int FLIP = 0>
GASMA_HD void scan_highway_loop(const uint32_t* base, int stride, int sc, int& sp, int& len) {
    constexpr int NW = 2 * W, BITS = 64 * W;
    sp = imax(sc, BITS);
    len = BITS;
    int w = sc >> 5;
    if (w >= NW) return;
    uint32_t prev = w > 0 ? base[(w - 1) * stride] : 0u;
    uint32_t cur = base[w * stride];
    uint32_t next = w + 1 < NW ? base[(w + 1) * stride] : 0u;
    uint32_t inv = 0xFFFFFFFFu, mask = 0xFFFFFFFFu << (sc & 31);
    for (;;) {
        const uint32_t t = (flip_word<FLIP>(prev, cur, next, w == 0, w == NW - 1) ^ inv) & mask;
        if (t) {
            const int pos = 32 * w + ctz32(t);
            if (inv == 0u) { len = pos - sp; return; }
            sp = pos;                                  // bit `pos` is 0, so it cannot match the search for a 1
            inv = 0u;
            mask = 0xFFFFFFFFu << (pos & 31);
        } else {
            if (++w >= NW) return;
            prev = cur;
            cur = next;
            next = w + 1 < NW ? base[(w + 1) * stride] : 0u;
            mask = 0xFFFFFFFFu;
        }
    }
}

// Continuation of a search whose highway start sp is known and whose flipped lane holds no hurdle in [sp, q):
// len = (first hurdle at/after bit q) - sp, BITS if none follows.  Words that start beyond the destination column de
// are not visited: a highway that long is clamped to the destination by the caller whatever its exact end.
template <int W, int FLIP = 0>
GASMA_HD int scan_highway_tail(const uint32_t* base, int stride, int q, int de, int sp) {
    constexpr int NW = 2 * W, BITS = 64 * W;
    int w = q >> 5;
    if (w >= NW) return BITS;
    uint32_t prev = w > 0 ? base[(w - 1) * stride] : 0u;
    uint32_t cur = base[w * stride];
    uint32_t mask = 0xFFFFFFFFu << (q & 31);
    for (;;) {
        const uint32_t next = w + 1 < NW ? base[(w + 1) * stride] : 0u;
        const uint32_t t = flip_word<FLIP>(prev, cur, next, w == 0, w == NW - 1) & mask;
        if (t) return 32 * w + ctz32(t) - sp;
        if (++w >= NW || 32 * w > de) return BITS;
        prev = cur;
        cur = next;
        mask = 0xFFFFFFFFu;
    }
}

GASMA_HD uint32_t funnel_r(uint32_t lo, uint32_t hi, int s) {  // low word of (hi:lo) >> s, 0 <= s < 32
#if defined(__CUDA_ARCH__)
    return __funnelshift_r(lo, hi, s);
#else
    return (uint32_t)((((uint64_t)hi << 32) | lo) >> s);
#endif
}

// Same result as scan_highway_loop.  Fast path: the 64 bits of the flipped lane that start at sc are gathered
// into one window and both searches are resolved without a branch; highways are geometrically distributed
// (~1/error-rate bases), so the loop is only needed when a run does not end inside the window.
// `de` is the lane's destination column: the caller clamps any highway with sp+len > de to the destination, so a
// run that leaves the window beyond de needs no exact end (len = BITS stands for "past the destination").
template <int W, int FLIP = 0>
GASMA_HD void scan_highway(const uint32_t* base, int stride, int sc, int de, int& sp, int& len) {
    constexpr int NW = 2 * W, BITS = 64 * W;
    const int w = sc >> 5, b = sc & 31;
    if (w >= NW) { sp = imax(sc, BITS); len = BITS; return; }
    const uint32_t om = w > 0 ? base[(w - 1) * stride] : 0u;
    const uint32_t o0 = base[w * stride];
    const uint32_t o1 = w + 1 < NW ? base[(w + 1) * stride] : 0u;
    const uint32_t o2 = w + 2 < NW ? base[(w + 2) * stride] : 0u;
    const uint32_t o3 = w + 3 < NW ? base[(w + 3) * stride] : 0u;
    // words beyond the vector are zero and must stay zero: they are never "last"
    const uint32_t f0 = flip_word<FLIP>(om, o0, o1, w == 0, w == NW - 1), f1 = flip_word<FLIP>(o0, o1, o2, false, w + 1 == NW - 1),
                   f2 = flip_word<FLIP>(o1, o2, o3, false, w + 2 == NW - 1);
    const uint32_t lo = funnel_r(f0, f1, b), hi = funnel_r(f1, f2, b);   // flipped lane bits [sc, sc+64), zero fill
    const uint32_t zlo = ~lo, zhi = ~hi;
    bool resolved = false;
    if ((zlo | zhi) != 0u) {
        const int fz = zlo ? ctz32(zlo) : 32 + ctz32(zhi);
        const uint32_t tlo = fz < 32 ? (lo & (0xFFFFFFFFu << fz)) : 0u;
        const uint32_t thi = fz < 32 ? hi : (hi & (0xFFFFFFFFu << (fz - 32)));
        if ((tlo | thi) != 0u) {
            sp = sc + fz;
            len = (tlo ? ctz32(tlo) : 32 + ctz32(thi)) - fz;
            resolved = true;
        } else {   // the highway leaves the window: no hurdle up to the end of the vector / past the destination, or go on
            sp = sc + fz;
            len = (sc + 64 >= BITS || sc + 64 > de) ? BITS : scan_highway_tail<W, FLIP>(base, stride, sc + 64, de, sp);
            resolved = true;
        }
    }
    if (!resolved) scan_highway_loop<W, FLIP>(base, stride, sc, sp, len);   // 64 hurdles in a row
}

// pop_count_between(from,to), utils.h:263-270 = shift_left(from).shift_right(from+BITS-to).pop_count():
// bits [from, to) clipped to the vector, 0 if from >= to, 0 if the second shift would be negative.  from >= 0.
template <int W>
GASMA_HD int pcb(const uint32_t* base, int stride, int from, int to) {
    constexpr int BITS = 64 * W;
    if (to > BITS + from) return 0;
    to = imin(to, BITS);
    if (from >= to) return 0;
    const int w0 = from >> 5, w1 = (to - 1) >> 5;
    const uint32_t mlo = 0xFFFFFFFFu << (from & 31), mhi = 0xFFFFFFFFu >> (31 - ((to - 1) & 31));
    const uint32_t a = base[w0 * stride] & mlo, z = base[w1 * stride] & mhi;   // w0 == w1: both masks on one word
    int c = w0 == w1 ? popc32(a & mhi) : popc32(a) + popc32(z);
    if (w1 - w0 >= 2) c += popc32(base[(w0 + 1) * stride]);   // up to three words without a loop
    // longer ranges are rare; the loop stays rolled (this function is inlined three times into a kernel whose speed
    // depends on its code staying inside the 32 KB instruction-cache tier)
#if defined(__CUDA_ARCH__)
#pragma unroll 1
#endif
    for (int w = w0 + 2; w < w1; ++w) c += popc32(base[w * stride]);
    return c;
}

template <int W>
GASMA_HD const uint32_t* lane_base(const LaneStore<W>& S, int li) { return S.lanes + (size_t)(li * 2 * W) * S.stride; }

// CIGAR ops of one pair while it is being aligned.  A pair's ops go to its PRIMARY ROW: kRowOps ops = one 32-byte
// sector, rows of consecutive pairs contiguous -- a 1 M-pair batch stages into 32 MB that stay in L2 and are read
// back coalesced.  The few pairs with more ops move to their OVERFLOW ROW (`cap` ops, enough for any traversal).
// Rule for consumers: a list of n <= kRowKeep ops is in the primary row, a longer one is complete in the overflow row.
//
// The kernel is bound by instruction issue and the three emit sites of _update_CIGAR run on every step of every
// warp, so emit() is an unconditional indexed store + increment.  Room for the (at most two) ops of the next
// update() is checked ONCE before it (reserve2): while n <= kRowKeep the primary row has it; beyond that the list is
// copied to the overflow row and continues there.
constexpr int kRowOps = 16;
constexpr int kRowKeep = kRowOps - 2;
#ifndef GASMA_ROW_PITCH
#define GASMA_ROW_PITCH 16
#endif
constexpr int kRowPitch = GASMA_ROW_PITCH;   // ops between the primary rows of consecutive pairs (>= kRowOps)
struct OpSink {
    uint16_t* rows;  // primary rows [pairs][kRowOps]; never null (a launch that wants no ops still stages them)
    uint16_t* ovf;   // overflow rows [pairs][cap]
    int cap;         // ops an overflow row holds
    uint32_t p;      // this pair's row index
    uint16_t* ops;   // row being written
    int n;
    GASMA_HD void begin(uint32_t pair) {
        p = pair;
        n = 0;
        ops = rows + (size_t)pair * kRowPitch;
    }
    GASMA_HD void emit(int code, int len) {
        ops[n] = (uint16_t)((len << 3) | code);
        ++n;
    }
    GASMA_HD void spill() {   // rare
        uint16_t* prim = rows + (size_t)p * kRowPitch;
        uint16_t* big = ovf + (size_t)p * cap;
        if (ops == prim) {
            for (int i = 0; i < n; ++i) big[i] = prim[i];
            ops = big;
        }
        // an overflow row that is full as well cannot happen (cap covers the longest traversal; the caller flags
        // n > cap); further ops then land on the primary row's first two slots instead of out of bounds
        if (n + 2 > cap) ops = prim - n;
    }
    GASMA_HD void reserve2() {
        if (n > kRowKeep) spill();
    }
    // _update_CIGAR, hurdle_matrix.h:238-251
    GASMA_HD void update(int best, int cur, int total) {
        reserve2();
        if (best < cur) emit(1, cur - best);
        else if (best > cur) emit(2, best - cur);
        if (total > 0) emit(0, total);
    }
    GASMA_HD void finish() { reserve2(); }   // a list that ended at kRowKeep+1 or +2 ops in the primary row moves over too
    // the list a consumer reads (n = the pair's op count)
    static GASMA_HD const uint16_t* list(const uint16_t* rows, const uint16_t* ovf_rows, int stride, long long p, uint32_t n) {
        return n <= (uint32_t)kRowKeep ? rows + (size_t)p * kRowPitch : ovf_rows + (size_t)p * stride;
    }
};

// Running position of one alignment (registers).
struct PairState {
    int m, n;        // clamped lengths
    int cl, cc;      // current lane / column
    int cost, steps;
    bool first;      // is_first_step (hurdle_matrix.h:572)
    int lb, ub;      // lane range of this pair; only read by the CORR instantiations (otherwise it is -k..k)
};

// Lane range of one pair (hurdle_matrix.h:634-644).  The reference's default build uses -k..k; built with
// -DCORRECTION it widens the band towards the destination lane n-m: [-k, n-m+k] for m <= n, [n-m-k, k] otherwise.
// CORR selects that variant at compile time, as the reference does, so the default kernels carry none of it.
template <bool CORR>
GASMA_HD void lane_range(int k, int m, int n, int& lb, int& ub) {
    lb = -k; ub = k;
    if (CORR) { if (m <= n) ub = n - m + k; else lb = n - m - k; }
}

// mask of the low n bits; n <= 0 gives 0, n >= 32 all ones
GASMA_HD uint32_t low_bits(int n) {
#if defined(__CUDA_ARCH__)
    return __funnelshift_lc(0xFFFFFFFFu, 0u, (uint32_t)imax(n, 0));   // clamps the shift at 32
#else
    return n <= 0 ? 0u : (n >= 32 ? 0xFFFFFFFFu : (0xFFFFFFFFu >> (32 - n)));
#endif
}

// 32-bit plane words of a string: word j = bits [32j, 32j+32) of the stream window that starts at bit `bit`, bits >= len
// cleared (the canonical 'A' tail), plus one zero word on top.  The stream must be readable for 2W+1 words from bit>>5.
template <int W>
GASMA_HD void extract_plane32(const uint32_t* stream32, int bit, int len, uint32_t out[2 * W + 1]) {
    constexpr int NW = 2 * W;
    const int idx = bit >> 5, sh = bit & 31;
    uint32_t prev = stream32[idx];
#pragma unroll
    for (int j = 0; j < NW; ++j) {
        const uint32_t nxt = stream32[idx + j + 1];
        const uint32_t v = funnel_r(prev, nxt, sh);
        out[j] = v & low_bits(len - 32 * j);
        prev = nxt;
    }
    out[NW] = 0u;
}

// reset(): highways::reset (:106-119) + _construct_hurdles (:441-455), straight from the 2-bit streams.
// Lane l < 0 compares (read >> |l|) with ref, lane l >= 0 compares (ref >> l) with read (a7).  The four planes are
// held as 32-bit words, so a shifted word is ONE funnel shift of two neighbours and a lane word is two more logic ops;
// shifts of 32 and more (bands beyond 31 lanes) re-read the shifted string from the stream instead.
// r0/r1 (f0/f1) = C|T and G|T plane streams of the read (ref), rbit/fbit = the string's first bit in them; the streams
// must be readable for 2W+3 words from the string's first word.
template <int W, bool CORR = false>
GASMA_HD void pair_init32(const KParams& P, const uint32_t* r0, const uint32_t* r1, int rbit, int m, const uint32_t* f0,
                          const uint32_t* f1, int fbit, int n, const LaneStore<W>& S, PairState& ps) {
    constexpr int BITS = 64 * W, NW = 2 * W;
    uint32_t A0[NW + 1], A1[NW + 1], B0[NW + 1], B1[NW + 1];
    extract_plane32<W>(r0, rbit, m, A0);
    extract_plane32<W>(r1, rbit, m, A1);
    extract_plane32<W>(f0, fbit, n, B0);
    extract_plane32<W>(f1, fbit, n, B1);
    int lb, ub;
    lane_range<CORR>(P.k, m, n, lb, ub);
    for (int l = lb; l <= ub; ++l) {
        const int a = iabs(l);
        uint32_t* b = S.lanes + (size_t)((l - lb) * NW) * S.stride;
        if (a >= 32) {   // rare: the shifted string comes from the stream again
            uint32_t T0[NW + 1], T1[NW + 1];
            extract_plane32<W>(l < 0 ? r0 : f0, (l < 0 ? rbit : fbit) + a, (l < 0 ? m : n) - a, T0);
            extract_plane32<W>(l < 0 ? r1 : f1, (l < 0 ? rbit : fbit) + a, (l < 0 ? m : n) - a, T1);
#pragma unroll
            for (int w = 0; w < NW; ++w)
                b[w * S.stride] = l < 0 ? ((T0[w] ^ B0[w]) | (T1[w] ^ B1[w])) : ((T0[w] ^ A0[w]) | (T1[w] ^ A1[w]));
        } else if (l < 0) {
#pragma unroll
            for (int w = 0; w < NW; ++w)
                b[w * S.stride] = (funnel_r(A0[w], A0[w + 1], a) ^ B0[w]) | (funnel_r(A1[w], A1[w + 1], a) ^ B1[w]);
        } else {
#pragma unroll
            for (int w = 0; w < NW; ++w)
                b[w * S.stride] = (funnel_r(B0[w], B0[w + 1], a) ^ A0[w]) | (funnel_r(B1[w], B1[w + 1], a) ^ A1[w]);
        }
        S.state[(l - lb) * S.stride] = LaneState<W>::pack(-1, 0, BITS, BITS);
    }
    ps.lb = lb; ps.ub = ub;
    ps.m = m; ps.n = n; ps.cl = 0; ps.cc = 0; ps.cost = 0; ps.steps = 0; ps.first = true;
}

// The same from already extracted 64-bit planes (host harness, callers that hold planes): the planes are laid out as
// zero-padded streams that start at bit 0.
template <int W, bool CORR = false>
GASMA_HD void pair_init(const KParams& P, const uint64_t A0[W], const uint64_t A1[W], const uint64_t B0[W],
                        const uint64_t B1[W], int m, int n, const LaneStore<W>& S, PairState& ps) {
    constexpr int NW = 2 * W;
    uint32_t s[4][NW + 4];
#pragma unroll
    for (int j = 0; j < NW + 4; ++j) {
        const int q = j >> 1;
        s[0][j] = j < NW ? (uint32_t)(A0[q] >> (32 * (j & 1))) : 0u;
        s[1][j] = j < NW ? (uint32_t)(A1[q] >> (32 * (j & 1))) : 0u;
        s[2][j] = j < NW ? (uint32_t)(B0[q] >> (32 * (j & 1))) : 0u;
        s[3][j] = j < NW ? (uint32_t)(B1[q] >> (32 * (j & 1))) : 0u;
    }
    pair_init32<W, CORR>(P, s[0], s[1], 0, m, s[2], s[3], 0, n, S, ps);
}

// One _step() (hurdle_matrix.h:407-434) including its _update_highway_list and _choose_best_highway.
// Returns true when the traversal is over (the caller then runs pair_finish).
template <int W, bool CORR = false, int FLIP = 0>
GASMA_HD bool pair_step(const KParams& P, const LaneStore<W>& S, PairState& ps, OpSink& sink, uint32_t& status) {
    constexpr int BITS = 64 * W;
    const int x = P.x, o = P.o, e = P.e;
    const int lb = CORR ? ps.lb : -P.k, ub = CORR ? ps.ub : P.k;
    const bool GLOBAL = P.global != 0;
    const int m = ps.m, n = ps.n, cl = ps.cl, cc = ps.cc;
    const bool leap_cost = GLOBAL || !ps.first;
    const int dl = n - m;
    if (++ps.steps > 4 * BITS + 16) { status |= kStatusSteps; return true; }

    // ---- _update_highway_list, loops 1 and 2 fused: both arg-max candidates are tracked, the one that
    // applies is picked once `reach` is known (hurdle_matrix.h:291-352) ----
    bool reach = false;
    using ST = typename LaneState<W>::type;
    // both maxima start below every candidate (the heuristics are finite), so the first lane is always taken
    double bestH_s = -1.0e300; int bestLeap_s = 0, bl_s = 0; ST st_s = 0;
    int bestH_r = INT32_MIN, bestLeap_r = 0, bl_r = 0; ST st_r = 0;
    for (int l = lb; l <= ub; ++l) {
        const int li = l - lb;
        const int sc = cc + sfc(cl, l);
        const int de = lane_destination(m, n, l);
        ST st = S.state[li * S.stride];
        int sp = st_sp(st), len = st_len(st), nsw = st_nsw(st);
        if (sp < sc) {  // cache miss: look up the closest highway of this lane (:293-310)
            nsw = iabs(l - cl);
            scan_highway<W, FLIP>(lane_base<W>(S, li), S.stride, sc, de, sp, len);
            if (sp + len > de) { len = imax(0, de - sp); reach = true; }
        }
        const int sw = leap_cost ? slp(cl, l, o, e) : 0;
        const int nh = pcb<W>(lane_base<W>(S, li), S.stride, sc, sp + len);
        const int hc = x * nh;
        st = LaneState<W>::pack(sp, len, nsw, nh);
        S.state[li * S.stride] = st;
        // significance heuristic (:328-332): ((ms*len) + (mms*nh)) + (ids*nsw), each op rounded
        const double h = dadd(dadd(dmul(P.ms, (double)len), dmul(P.mms, (double)nh)), dmul(P.ids, (double)nsw));
        const int lh = -sw;
        if (h > bestH_s || (h == bestH_s && lh > bestLeap_s)) {
            bestH_s = h; bestLeap_s = lh; bl_s = l; st_s = st;
        }
        // destination heuristic (:334-342), integer valued
        const int fsw = GLOBAL ? slp(l, dl, o, e) : 0;
        const int hr = -sw - hc - fsw - x * (de - sp - len);
        const int lhr = -sw - fsw;
        if (hr > bestH_r || (hr == bestH_r && lhr > bestLeap_r)) {
            bestH_r = hr; bestLeap_r = lhr; bl_r = l; st_r = st;
        }
    }
    const int bl = reach ? bl_r : bl_s;
    const ST st_bl = reach ? st_r : st_s;
    if (st_len(st_bl) <= 0) return true;  // no usable highway (:358-360, :570)

    // ---- _choose_best_highway (:368-401) ----
    const int spb = st_sp(st_bl);
    const int swb = leap_cost ? slp(cl, bl, o, e) : 0;
    int sT = x * st_nh(st_bl) + swb, sI = sT, b = bl;
    ST st_b = st_bl;
    const uint32_t* ob = lane_base<W>(S, bl - lb);
    for (int l = lb; l <= ub; ++l) {
        if (l == bl) continue;
        const ST st = S.state[(l - lb) * S.stride];
        const int sp = st_sp(st), s_l = sfc(l, bl);
        if (sp + s_l > spb) continue;
        const int end = sp + st_len(st);
        const int swl = leap_cost ? slp(cl, l, o, e) : 0;
        const int ic = swl + st_nh(st);  // == switch_cost + pop_count_between(cc+sfc(cl,l), end), no *x (:388)
        const int hp = x * pcb<W>(ob, S.stride, s_l + end, spb);
        const int tc = ic + slp(l, bl, o, e) + imax(0, hp);
        if (tc <= sT && ic <= sI) { sT = tc; sI = ic; b = l; st_b = st; }
    }
    // ---- _step (:412-433) ----
    const int swc = leap_cost ? slp(cl, b, o, e) : 0;
    ps.cost += swc + x * st_nh(st_b);
    const int endb = st_sp(st_b) + st_len(st_b);
    sink.update(b, cl, endb - (cc + sfc(cl, b)));
    ps.cl = b;
    ps.cc = endb;
    ps.first = false;
    return endb >= lane_destination(m, n, b);
}

// run() fix-up (hurdle_matrix.h:575-590).  A destination lane outside the band keeps the state of a fresh
// object: lane value 1 (utils.h:57-59) and destination -|lane| (hurdle_matrix.h:79-92 with m=n=0).
template <int W, bool CORR = false>
GASMA_HD void pair_finish(const KParams& P, const LaneStore<W>& S, PairState& ps, OpSink& sink) {
    const int x = P.x;
    const int lb = CORR ? ps.lb : -P.k, ub = CORR ? ps.ub : P.k;
    const int dl = ps.n - ps.m, cl = ps.cl, cc = ps.cc;
    const bool in_band = dl >= lb && dl <= ub;   // always true with CORR
    const int dc = in_band ? lane_destination(ps.m, ps.n, dl) : -iabs(dl);
    if (cl != dl || cc < dc) {
        int d = 0;
        if (in_band) d = pcb<W>(lane_base<W>(S, dl - lb), S.stride, cc + sfc(cl, dl), dc);
        // out of band: dc < 0 <= from, so pop_count_between of the default value 1 is 0 as well
        ps.cost += (P.global ? slp(cl, dl, P.o, P.e) : 0) + imax(0, x * d);
        sink.update(dl, cl, d);
    }
}

// The whole alignment of one pair (host harness, and kernels that do not interleave pairs).  Returns status bits.
template <int W, bool CORR = false, int FLIP = 0>
GASMA_HD uint32_t greedy_pair(const KParams& P, const uint64_t A0[W], const uint64_t A1[W], const uint64_t B0[W],
                              const uint64_t B1[W], int m, int n, const LaneStore<W>& S, OpSink& sink, int& cost_out,
                              int& steps_out) {
    PairState ps;
    uint32_t status = 0;
    pair_init<W, CORR>(P, A0, A1, B0, B1, m, n, S, ps);
    while (!pair_step<W, CORR, FLIP>(P, S, ps, sink, status)) {}
    pair_finish<W, CORR>(P, S, ps, sink);
    cost_out = ps.cost;
    steps_out = ps.steps;
    return status;
}

}  // namespace gasma
