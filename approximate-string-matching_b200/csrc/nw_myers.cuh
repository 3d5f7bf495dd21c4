// nw_myers.cuh -- K4 fast path: the NW verifier for unit penalties (x = o = e = 1), bit-parallel.
//
// The reference's benchmark always verifies with benchmark(1,1,1,...) (GASMA/benchmark/benchmark.cpp:22,27),
// i.e. parasail_nw_trace with match 0 / mismatch -1 / gap -(1+(g-1)) (benchmark_utils.h:113-116,139-142,288):
// plain Levenshtein distance.  For that scoring the DP matrix is computed one COLUMN per step with Myers' /
// Hyyro's bit-vector recurrence on the vertical and horizontal deltas (all rows of the read at once, 32 rows per
// machine word) instead of one cell at a time, in its global (NW) form: every column enters row 0 with a
// horizontal delta of +1 (H[0][j] = j) and column 0 has vertical deltas +1 (H[i][0] = i).
//
// Traceback is the same deterministic rule as the general Gotoh kernel / oracle (oracle/gasma_oracle.c:go_nw;
// with o == e a gap is never "extended" in preference to being re-opened, so the three-state walk collapses):
//   at (i,j): diagonal if H[i][j] == H[i-1][j-1] + (read[i] != ref[j]), else I (up) if H[i][j] == H[i-1][j] + 1,
//   else D (left).
// Both tests are bits of two vectors that fall out of the column step, so two bit-vectors per column are all the
// traceback needs:  UP = Pv (vertical delta +1)  and
//   DG = Eq | ~zero(dv + dh'),   dv = vertical delta of column j at row i, dh' = horizontal delta at row i-1
// because H[i][j] - H[i-1][j-1] = dv_j(i) + dh_j(i-1) is 0 or 1, is 0 on a match, and a mismatching diagonal is
// optimal exactly when that difference is 1.
//
// Two forms live here: full columns (2W words, penalty only: myers_forward, used when no upper bound on the
// optimum is known) and the banded form (one 32-bit word per column, penalty + traceback in one pass, below).
// Bands wider than 32 diagonals go to the general Gotoh kernel (same traceback rule).
#pragma once
#include "common.cuh"
#include "nw.cuh"

namespace gasma {

// One column.  NW32 = number of 32-bit words covering the read.  eq[] = rows whose character equals ref[j].
// Updates pv/mv in place, returns the horizontal delta at row `m` (bit m-1), writes DG into dg[].
template <int NW32>
GASMA_HD int myers_column(const uint32_t eq[NW32], uint32_t pv[NW32], uint32_t mv[NW32], uint32_t dg[NW32], int m) {
    uint32_t ph[NW32], mh[NW32], xv[NW32];
    uint32_t carry = 0u;
#pragma unroll
    for (int w = 0; w < NW32; ++w) {
        xv[w] = eq[w] | mv[w];
        // sum = (eq & pv) + pv over the whole multi-word vector
        const uint32_t a = eq[w] & pv[w];
        const uint64_t s = (uint64_t)a + pv[w] + carry;
        carry = (uint32_t)(s >> 32);
        const uint32_t xh = ((uint32_t)s ^ pv[w]) | eq[w];
        ph[w] = mv[w] | ~(xh | pv[w]);
        mh[w] = pv[w] & xh;
    }
    int dh = 0;
    if (m > 0) {
        const int w = (m - 1) >> 5, b = (m - 1) & 31;
        uint32_t p = 0u, q = 0u;
#pragma unroll
        for (int j = 0; j < NW32; ++j)
            if (j == w) { p = ph[j]; q = mh[j]; }
        dh = (int)((p >> b) & 1u) - (int)((q >> b) & 1u);
    }
    // shift the horizontal deltas down one row; row 0 enters with +1 (global alignment)
    uint32_t cp = 1u, cm = 0u;
#pragma unroll
    for (int w = 0; w < NW32; ++w) {
        const uint32_t php = (ph[w] << 1) | cp, mhp = (mh[w] << 1) | cm;
        cp = ph[w] >> 31;
        cm = mh[w] >> 31;
        const uint32_t npv = mhp | ~(xv[w] | php);
        const uint32_t nmv = php & xv[w];
        const uint32_t zero = (~(npv | nmv) & ~(php | mhp)) | (npv & mhp) | (nmv & php);   // dv + dh' == 0
        dg[w] = eq[w] | ~zero;
        pv[w] = npv;
        mv[w] = nmv;
    }
    return dh;
}

GASMA_HD uint32_t funnel32(uint32_t lo, uint32_t hi, int s) {  // low word of (hi:lo) >> s, 0 <= s < 32
#if defined(__CUDA_ARCH__)
    return __funnelshift_r(lo, hi, s);
#else
    return (uint32_t)((((uint64_t)hi << 32) | lo) >> s);
#endif
}

// first row (0-based bit index) of the window kept for column j0 (0-based) when the band's top diagonal is hi
GASMA_HD int myers_window_start(int j0, int hi) { return imax(0, j0 - hi); }

// Penalty-only forward pass: H[m][n].
template <int W>
GASMA_HD int myers_forward(const uint64_t A0[W], const uint64_t A1[W], const uint64_t B0[W], const uint64_t B1[W], int m, int n) {
    constexpr int NW32 = 2 * W;
    if (m == 0) return n;
    uint32_t a0[NW32], a1[NW32], pv[NW32], mv[NW32];
#pragma unroll
    for (int j = 0; j < W; ++j) {
        a0[2 * j] = (uint32_t)A0[j]; a0[2 * j + 1] = (uint32_t)(A0[j] >> 32);
        a1[2 * j] = (uint32_t)A1[j]; a1[2 * j + 1] = (uint32_t)(A1[j] >> 32);
    }
#pragma unroll
    for (int w = 0; w < NW32; ++w) { pv[w] = 0xFFFFFFFFu; mv[w] = 0u; }
    int score = m;
    uint64_t b0 = 0, b1 = 0;
    for (int j = 0; j < n; ++j) {
        if ((j & 63) == 0) {   // next 64 reference characters
#pragma unroll
            for (int q = 0; q < W; ++q)
                if (q == (j >> 6)) { b0 = B0[q]; b1 = B1[q]; }
        }
        const uint32_t k0 = (uint32_t)0 - (uint32_t)(b0 & 1ull), k1 = (uint32_t)0 - (uint32_t)(b1 & 1ull);
        b0 >>= 1;
        b1 >>= 1;
        uint32_t eq[NW32], dg[NW32];
#pragma unroll
        for (int w = 0; w < NW32; ++w) eq[w] = ~(a0[w] ^ k0) & ~(a1[w] ^ k1);
        score += myers_column<NW32>(eq, pv, mv, dg, m);   // dg is dead here and compiled away
    }
    return score;
}

// ---- banded variant: the whole column is ONE 32-bit word that slides down the diagonal ---------------------
// With an upper bound on the optimum (the greedy penalty, or the exact optimum of a penalty-only pass) every optimal
// path stays on the diagonals d = j - i of nw_band(bound) (nw.cuh).  When those fit a 32-bit word, column j keeps only
// the rows i = j - hi + b, b = 0..31 (hi = the band's top diagonal): the window moves one row down per column, so
// between columns the vertical deltas shift by one bit.  The recurrence is the ordinary column step on that row set
// with the two cells the band cuts off taken as "+1 from their in-band neighbour":
//   * the cell above the window (diagonal hi+1) = its left neighbour + 1   -> horizontal delta +1 enters at bit 0
//   * the cell left of the bottom row (diagonal hi-32) = the cell above it + 1 -> vertical delta +1 enters at bit 31
// Both are upper bounds of the true values that a diagonal step always matches or beats, so every in-band cell holds
// the cost of the best path INSIDE the band (>= the true value, == on every optimal path), and the traceback tests
// "diagonal / up is optimal" come out exactly as on the full matrix for every cell an optimal walk visits.
// Rows i <= 0 (window still above the matrix, columns j < hi) are virtual: H(i,j) = j + |i|, i.e. vertical deltas
// of -1, which the recurrence keeps by itself and which makes row 0 read H(0,j) = j.  Rows i > m compare against
// the 'A' tail; nothing above them depends on them.
// Cost per column is independent of the read length (one word instead of 2W), and the vectors the traceback wants
// ARE the windows, so penalty and traceback come out of one pass.
//
// Traceback storage: one 8-byte entry per column in shared memory, tb[slot * stride]:
//   .dg : DG                      (diagonal step optimal)
//   .sel: DG ? Eq : UP            (match vs mismatch on a diagonal step, insert vs delete otherwise)
struct alignas(8) TraceCol { uint32_t dg, sel; };
template <int W>
GASMA_HD uint32_t word32_of(const uint64_t v[W], int w) {   // 32-bit word w of the vector; 0 outside
    uint64_t x = 0;
#pragma unroll
    for (int j = 0; j < W; ++j)
        if (j == (w >> 1)) x = v[j];
    return (uint32_t)(x >> ((w & 1) * 32));
}
GASMA_HD uint32_t funnel32_l(uint32_t lo, uint32_t hi, int s) {  // high word of (hi:lo) << s, 0 <= s < 32
#if defined(__CUDA_ARCH__)
    return __funnelshift_l(lo, hi, s);
#else
    return (uint32_t)(((((uint64_t)hi << 32) | lo) << s) >> 32);
#endif
}
GASMA_HD uint32_t brev32(uint32_t x) {
#if defined(__CUDA_ARCH__)
    return __brev(x);
#else
    x = ((x >> 1) & 0x55555555u) | ((x & 0x55555555u) << 1);
    x = ((x >> 2) & 0x33333333u) | ((x & 0x33333333u) << 2);
    x = ((x >> 4) & 0x0F0F0F0Fu) | ((x & 0x0F0F0F0Fu) << 4);
    x = ((x >> 8) & 0x00FF00FFu) | ((x & 0x00FF00FFu) << 8);
    return (x >> 16) | (x << 16);
#endif
}

// does the band of `bound` fit the 32-bit window?  (hi - 31 <= lo: the window then covers the whole band)
GASMA_HD bool myers_band_fits(const NwBand& bd) { return bd.hi - bd.lo + 1 <= 32; }

// Columns [1, j_end] of the banded matrix (1-based; j_end <= n).  When tb != nullptr the windows of the columns
// j > j_store are kept at slot j - 1 - j_store.  Returns H[m][j_end] (meaningful when row m is inside the last
// window, which it is for j_end == n).
template <int W>
GASMA_HD int myers_band_forward(const uint64_t A0[W], const uint64_t A1[W], const uint64_t B0[W], const uint64_t B1[W],
                                int m, int hi, int j_store, int j_end, TraceCol* tb, int stride) {
    // column 0: H(i,0) = |i|  ->  +1 below row 0, -1 at and above it; bit b is row b - hi
    uint32_t pv = hi >= 31 ? 0u : 0xFFFFFFFFu << (hi + 1);
    uint32_t mv = ~pv;
    int diag0 = 0;   // zero-cost diagonal steps of the window's top cell
    // read planes shifted up by hi, so that the window of (0-based) column c is bits [c, c+32) of the shifted plane
    uint32_t c0 = word32_of<W>(A0, 0), c1 = word32_of<W>(A1, 0);
    uint32_t wl0 = funnel32_l(0u, c0, hi), wl1 = funnel32_l(0u, c1, hi);
    TraceCol* out = tb;
    for (int cb = 0; 32 * cb < j_end; ++cb) {
        const uint32_t n0 = word32_of<W>(A0, cb + 1), n1 = word32_of<W>(A1, cb + 1);
        const uint32_t wh0 = funnel32_l(c0, n0, hi), wh1 = funnel32_l(c1, n1, hi);
        c0 = n0; c1 = n1;
        uint32_t r0 = brev32(word32_of<W>(B0, cb)), r1 = brev32(word32_of<W>(B1, cb));   // next ref character in bit 31
        const int cend = imin(32, j_end - 32 * cb);
        for (int s = 0; s < cend; ++s) {
            const uint32_t k0 = (uint32_t)((int32_t)r0 >> 31), k1 = (uint32_t)((int32_t)r1 >> 31);
            r0 <<= 1;
            r1 <<= 1;
            const uint32_t eq = ~(funnel32(wl0, wh0, s) ^ k0) & ~(funnel32(wl1, wh1, s) ^ k1);
            const uint32_t pvi = (pv >> 1) | 0x80000000u, mvi = mv >> 1;   // slide down one row
            const uint32_t xv = eq | mvi;
            const uint32_t xh = ((((eq & pvi) + pvi) ^ pvi)) | eq;
            const uint32_t ph = mvi | ~(xh | pvi), mh = pvi & xh;
            const uint32_t php = (ph << 1) | 1u, mhp = mh << 1;
            pv = mhp | ~(xv | php);
            mv = php & xv;
            const uint32_t zero = xh | mvi;   // Hyyro's D0: H(i,j) == H(i-1,j-1), i.e. dv + dh' == 0
            diag0 += (int)(zero & 1u);
            if (tb && 32 * cb + s >= j_store) {
                const uint32_t dg = eq | ~zero;
                *out = TraceCol{dg, eq | (~dg & pv)};
                out += stride;
            }
        }
        wl0 = wh0; wl1 = wh1;
    }
    // top cell of the last window: H(-hi,0) = hi, +1 per column unless the diagonal step was free; then down to row m
    const int b = imax(-1, m - j_end + hi);   // < 0 only in a partial recomputation, whose result is not used
    const uint32_t mask = (b >= 31 ? 0xFFFFFFFFu : ((1u << (b + 1)) - 1u)) & ~1u;
    int score = hi + j_end - diag0;
#if defined(__CUDA_ARCH__)
    score += __popc(pv & mask) - __popc(mv & mask);
#else
    score += __builtin_popcount(pv & mask) - __builtin_popcount(mv & mask);
#endif
    return score;
}

// Resumable walk back: emits runs while j > j_lo (and i > 0), reading the entries stored with base column j_store.
// A finished run goes to the sink: sink.flush(code, length, i) with i = the read row the run started after (walking
// backwards), i.e. the run covers the 0-based read positions [i, i + rows it consumed).
struct MyersTrace {
    int i, j, b;      // b = i - j + hi: the bit of cell (i, j) in column j's window
    int cur, run;     // the run being extended
};
// Sink 1: CIGAR ops, written backwards from the end of the pair's staging row (the walk runs from the end of the
// alignment); the row holds 2*max_length+8 ops, more than any walk can emit, so there is no capacity test.
struct OpsBackSink {
    uint16_t* wp;     // next op is written at *--wp
    GASMA_HD void flush(int code, int run, int) { *--wp = (uint16_t)((run << 3) | code); }
};
template <class Sink>
GASMA_HD void myers_trace_push(MyersTrace& t, Sink& sink, int code, int len) {
    if (code != t.cur) {
        if (t.cur >= 0) sink.flush(t.cur, t.run, t.i);
        t.cur = code;
        t.run = 0;
    }
    t.run += len;
}
template <class Sink>
GASMA_HD void myers_trace_segment(MyersTrace& t, Sink& sink, int j_lo, int j_store, const TraceCol* tb, int stride) {
    const TraceCol* col = tb + (size_t)(t.j - 1 - j_store) * stride;
    while (t.i > 0 && t.j > j_lo) {
        const TraceCol c = *col;
        const int dg = (int)((c.dg >> (t.b & 31)) & 1u), sel = (int)((c.sel >> (t.b & 31)) & 1u);
        // dg: '=' (3) / 'X' (4) by sel;  otherwise I (1) when sel (up is optimal), else D (2)
        myers_trace_push(t, sink, dg ? 4 - sel : 2 - sel, 1);
        const int di = dg | sel, dj = dg | (sel ^ 1);
        t.i -= di;
        t.j -= dj;
        t.b += dj - di;
        col -= dj ? stride : 0;
    }
}
// the boundary tail (first row / first column) and the flush of the last run
template <class Sink>
GASMA_HD void myers_trace_finish(MyersTrace& t, Sink& sink) {
    if (t.i > 0) { myers_trace_push(t, sink, 1, t.i); t.i = 0; }
    if (t.j > 0) myers_trace_push(t, sink, 2, t.j);
    myers_trace_push(t, sink, -2, 0);
}

// Penalty and (when tb != nullptr) traceback of one alignment whose band fits the window.  Only 1/kTraceParts of the
// columns is resident at a time: the last part is stored and walked first, then the columns below are recomputed
// part by part into the same slots ((kTraceParts+1)/2 x the column work for 1/kTraceParts of the shared memory,
// i.e. that many more alignments in flight per SM).  tb holds myers_trace_slots(n) entries per alignment.
// Returns the penalty; the runs of the walk go to `sink`.
#ifndef GASMA_TRACE_PARTS
#define GASMA_TRACE_PARTS 3
#endif
constexpr int kTraceParts = GASMA_TRACE_PARTS;
GASMA_HD int myers_part_start(int n, int part) { return (int)((long long)n * part / kTraceParts); }
GASMA_HD constexpr int myers_trace_slots(int max_n) { return (max_n + kTraceParts - 1) / kTraceParts; }
template <int W, class Sink>
GASMA_HD int myers_band_align(const uint64_t A0[W], const uint64_t A1[W], const uint64_t B0[W], const uint64_t B1[W],
                              int m, int n, int hi, TraceCol* tb, int stride, Sink& sink) {
    if (m == 0 || n == 0) {
        if (tb) { MyersTrace t{m, n, 0, -1, 0}; myers_trace_finish(t, sink); }
        return imax(m, n);
    }
    if (!tb) return myers_band_forward<W>(A0, A1, B0, B1, m, hi, 0, n, nullptr, 0);
    MyersTrace t{m, n, m - n + hi, -1, 0};
    int pen = 0;
#pragma unroll
    for (int part = kTraceParts - 1; part >= 0; --part) {
        if (!(t.i > 0 && t.j > 0)) break;
        const int lo = myers_part_start(n, part);
        const int h = myers_band_forward<W>(A0, A1, B0, B1, m, hi, lo, t.j, tb, stride);   // t.j == n the first time
        if (part == kTraceParts - 1) pen = h;
        myers_trace_segment(t, sink, lo, lo, tb, stride);
    }
    myers_trace_finish(t, sink);
    return pen;
}

// ---- the walk without a CIGAR: what the coverage counter needs from it, as bit masks over read positions ----
// benchmark_coverage.h:26-67 with threshold 3 keeps the read positions under '=' runs of length >= 3.  The walk
// consumes the read rows in descending order, so it just records, per consumed row p (0-based position),
//   E[p] = the row was consumed by an '=' step,   B[p] = a D step happened between row p and the row above it
// (a D does not consume a row but ends a run), 32 positions per word, accumulated in registers and flushed when a
// word is complete.  No run lengths, no per-run branches; metrics.cuh:runs3_positions turns E, B into the positions.
template <int W>
struct CoverTrace {
    int i, j, b;
    uint32_t e, brk;       // word under construction
    uint32_t pend;         // a D step since the last consumed row
    uint32_t E[2 * W], B[2 * W];
};
template <int W>
GASMA_HD void myers_cover_segment(CoverTrace<W>& t, int j_lo, int j_store, const TraceCol* tb, int stride) {
    const TraceCol* col = tb + (size_t)(t.j - 1 - j_store) * stride;
    while (t.i > 0 && t.j > j_lo) {
        const TraceCol c = *col;
        const uint32_t dg = (c.dg >> (t.b & 31)) & 1u, sel = (c.sel >> (t.b & 31)) & 1u;
        const uint32_t di = dg | sel, dj = dg | (sel ^ 1u);   // diagonal: both; up (I): row only; left (D): column only
        const int p = t.i - 1;
        const uint32_t bit = di << (p & 31);
        t.e |= (dg & sel) ? bit : 0u;
        t.brk |= t.pend ? bit : 0u;
        t.pend = di ^ 1u;
        if (di && (p & 31) == 0) {
            t.E[p >> 5] = t.e;
            t.B[p >> 5] = t.brk;
            t.e = t.brk = 0u;
        }
        t.i -= (int)di;
        t.j -= (int)dj;
        t.b += (int)dj - (int)di;
        col -= dj ? stride : 0;
    }
}
// Penalty + E/B masks of one alignment whose band fits the window (same part-resident scheme as myers_band_align).
template <int W>
GASMA_HD int myers_band_cover(const uint64_t A0[W], const uint64_t A1[W], const uint64_t B0[W], const uint64_t B1[W],
                              int m, int n, int hi, TraceCol* tb, int stride, CoverTrace<W>& t) {
    t.i = m; t.j = n; t.b = m - n + hi;
    t.e = t.brk = t.pend = 0u;
#pragma unroll
    for (int w = 0; w < 2 * W; ++w) { t.E[w] = 0u; t.B[w] = 0u; }
    if (m == 0 || n == 0) return imax(m, n);   // only I or only D: no '=' at all
    int pen = 0;
#pragma unroll
    for (int part = kTraceParts - 1; part >= 0; --part) {
        if (!(t.i > 0 && t.j > 0)) break;
        const int lo = myers_part_start(n, part);
        const int h = myers_band_forward<W>(A0, A1, B0, B1, m, hi, lo, t.j, tb, stride);
        if (part == kTraceParts - 1) pen = h;
        myers_cover_segment<W>(t, lo, lo, tb, stride);
    }
    if (t.i > 0) {   // rows left at column 0 are an I run; the word under construction still has to land
        t.E[(t.i - 1) >> 5] = t.e;
        t.B[(t.i - 1) >> 5] = t.brk;
    }
    return pen;
}

}  // namespace gasma
