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
    // 64 reference characters per block; the block loop is unrolled so that B0[q] / B1[q] have compile-time indices
    // (a run-time index would put every plane array of the calling kernel into local memory)
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
    for (int q = 0; q < W; ++q) {
        if (64 * q >= n) continue;
        uint64_t b0 = B0[q], b1 = B1[q];
        const int jend = imin(64, n - 64 * q);
#if defined(__CUDA_ARCH__)
#pragma unroll 1
#endif
        for (int j = 0; j < jend; ++j) {
            const uint32_t k0 = (uint32_t)0 - (uint32_t)(b0 & 1ull), k1 = (uint32_t)0 - (uint32_t)(b1 & 1ull);
            b0 >>= 1;
            b1 >>= 1;
            uint32_t eq[NW32], dg[NW32];
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
            for (int w = 0; w < NW32; ++w) eq[w] = ~(a0[w] ^ k0) & ~(a1[w] ^ k1);
            score += myers_column<NW32>(eq, pv, mv, dg, m);   // dg is dead here and compiled away
        }
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
// (x >> 1) | 0x80000000 and (x << 1) | 1 as ONE funnel shift each (the verifier's column loop is bound by the ALU pipe)
GASMA_HD uint32_t shr1_fill(uint32_t x) {
#if defined(__CUDA_ARCH__)
    return __funnelshift_r(x, 0xFFFFFFFFu, 1);
#else
    return (x >> 1) | 0x80000000u;
#endif
}
GASMA_HD uint32_t shl1_fill(uint32_t x) {
#if defined(__CUDA_ARCH__)
    return __funnelshift_l(0xFFFFFFFFu, x, 1);
#else
    return (x << 1) | 1u;
#endif
}
// 32-bit word w of the vector for a COMPILE-TIME w (callers unroll): no run-time index into the register array
template <int W>
GASMA_HD uint32_t half32(const uint64_t v[W], int w) { return (uint32_t)(v[w >> 1] >> ((w & 1) * 32)); }
GASMA_HD uint32_t funnel32_l(uint32_t lo, uint32_t hi, int s) {  // high word of (hi:lo) << s, 0 <= s < 32
#if defined(__CUDA_ARCH__)
    return __funnelshift_l(lo, hi, s);
#else
    return (uint32_t)(((((uint64_t)hi << 32) | lo) << s) >> 32);
#endif
}
GASMA_HD int popc32m(uint32_t x) {
#if defined(__CUDA_ARCH__)
    return __popc(x);
#else
    return __builtin_popcount(x);
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

// State of the banded recurrence between two columns: the vertical deltas of the window (pv / mv) and the number of
// free diagonal steps its top cell has taken so far.  Column 0: H(i,0) = |i| -> +1 below row 0, -1 at and above it;
// bit b of the window of column j is row j - hi + b.
struct BandCols {
    uint32_t pv, mv;
    int diag0;
};
GASMA_HD BandCols myers_band_start(int hi) {
    const uint32_t pv = hi >= 31 ? 0u : 0xFFFFFFFFu << (hi + 1);
    return BandCols{pv, ~pv, 0};
}

#ifndef GASMA_NW_COL_UNROLL
#define GASMA_NW_COL_UNROLL 1   // measured 1 / 2 / 4: 0.587 / 0.595 / 0.592 ms per 1 M pairs (the loop is ALU-bound, not overhead-bound)
#endif
constexpr int kNwColUnroll = GASMA_NW_COL_UNROLL;   // unroll factor of the column loop (loop overhead vs code size)

// Columns j_begin+1 .. j_end (1-based) of the banded matrix, continuing from the state `c` left after column j_begin
// (myers_band_start for j_begin == 0).  When tb != nullptr the windows of these columns are kept at slots 0, 1, ...
// (slot s = column j_begin + 1 + s).  Resumable at any column, so a caller can checkpoint `c` and recompute a range
// of columns later without starting from column 0.
template <int W>
GASMA_HD void myers_band_run(const uint64_t A0[W], const uint64_t A1[W], const uint64_t B0[W], const uint64_t B1[W], int hi,
                             int j_begin, int j_end, BandCols& c, TraceCol* tb, int stride) {
    uint32_t pv = c.pv, mv = c.mv;
    int diag0 = c.diag0;
    TraceCol* out = tb;
    // One 32-column block at a time.  The loop over blocks is unrolled with a range test per block so that every plane
    // word is named by a compile-time index: a run-time index would put the plane arrays into local memory.
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
    for (int cb = 0; cb < 2 * W; ++cb) {
        if (32 * cb >= j_end || 32 * (cb + 1) <= j_begin) continue;
        const int s0 = imax(j_begin - 32 * cb, 0), cend = imin(32, j_end - 32 * cb);
        // read planes shifted up by hi, so that the window of (0-based) column q is bits [q, q+32) of the shifted plane
        const uint32_t al0 = cb > 0 ? half32<W>(A0, cb - 1) : 0u, am0 = half32<W>(A0, cb), ah0 = cb + 1 < 2 * W ? half32<W>(A0, cb + 1) : 0u;
        const uint32_t al1 = cb > 0 ? half32<W>(A1, cb - 1) : 0u, am1 = half32<W>(A1, cb), ah1 = cb + 1 < 2 * W ? half32<W>(A1, cb + 1) : 0u;
        const uint32_t wl0 = funnel32_l(al0, am0, hi), wh0 = funnel32_l(am0, ah0, hi);
        const uint32_t wl1 = funnel32_l(al1, am1, hi), wh1 = funnel32_l(am1, ah1, hi);
        uint32_t r0 = brev32(half32<W>(B0, cb)) << s0, r1 = brev32(half32<W>(B1, cb)) << s0;   // next ref character in bit 31
        uint32_t dsr = 0u;   // bit 0 of every column's D0 word, shifted in from the top: one instruction per column
#if defined(__CUDA_ARCH__)
#pragma unroll(kNwColUnroll)
#endif
        for (int s = s0; s < cend; ++s) {
            const uint32_t k0 = (uint32_t)((int32_t)r0 >> 31), k1 = (uint32_t)((int32_t)r1 >> 31);
            r0 <<= 1;
            r1 <<= 1;
            const uint32_t eq = ~(funnel32(wl0, wh0, s) ^ k0) & ~(funnel32(wl1, wh1, s) ^ k1);
            const uint32_t pvi = shr1_fill(pv), mvi = mv >> 1;   // slide down one row: +1 enters at the bottom
            const uint32_t xv = eq | mvi;
            const uint32_t xh = ((((eq & pvi) + pvi) ^ pvi)) | eq;
            const uint32_t ph = mvi | ~(xh | pvi), mh = pvi & xh;
            const uint32_t php = shl1_fill(ph), mhp = mh << 1;   // +1 enters at the top
            pv = mhp | ~(xv | php);
            mv = php & xv;
            const uint32_t zero = xh | mvi;   // Hyyro's D0: H(i,j) == H(i-1,j-1), i.e. dv + dh' == 0
            dsr = funnel32(dsr, zero, 1);      // (dsr >> 1) | (zero << 31): the top cell's free diagonal steps
            if (tb) {
                const uint32_t dg = eq | ~zero;
                *out = TraceCol{dg, eq | (~dg & pv)};
                out += stride;
            }
        }
        diag0 += popc32m(dsr);   // at most 32 columns per block, so no collected bit has been shifted out
    }
    c.pv = pv; c.mv = mv; c.diag0 = diag0;
}

// H[m][j] from the state after column j (row m must be inside that column's window, which it is for j == n):
// top cell of the window: H(-hi,0) = hi, +1 per column unless the diagonal step was free; then down to row m.
GASMA_HD int myers_band_score(const BandCols& c, int m, int j, int hi) {
    const int b = imax(-1, m - j + hi);
    const uint32_t mask = (b >= 31 ? 0xFFFFFFFFu : ((1u << (b + 1)) - 1u)) & ~1u;
    return hi + j - c.diag0 + popc32m(c.pv & mask) - popc32m(c.mv & mask);
}

// Columns [1, j_end] from column 0 (penalty-only callers).  Returns H[m][j_end].
template <int W>
GASMA_HD int myers_band_forward(const uint64_t A0[W], const uint64_t A1[W], const uint64_t B0[W], const uint64_t B1[W],
                                int m, int hi, int j_end) {
    BandCols c = myers_band_start(hi);
    myers_band_run<W>(A0, A1, B0, B1, hi, 0, j_end, c, nullptr, 0);
    return myers_band_score(c, m, j_end, hi);
}

// Resumable walk back: emits runs while j > j_lo (and i > 0), reading the entries of the columns j_lo+1 .. stored at
// slots 0, 1, ... (slot = column - 1 - j_lo).
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
// a walker bound to its sink: the part driver below only needs live / col / start / segment / finish
template <class Sink>
struct OpsWalker {
    MyersTrace t;
    Sink& sink;
    GASMA_HD bool live() const { return t.i > 0 && t.j > 0; }
    GASMA_HD int col() const { return t.j; }
    GASMA_HD void start(int m, int n, int hi) { t = MyersTrace{m, n, m - n + hi, -1, 0}; }
    GASMA_HD void segment(int j_lo, const TraceCol* tb, int stride) {
        const TraceCol* col = tb + (size_t)(t.j - 1 - j_lo) * stride;
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
    GASMA_HD void finish() {
        if (t.i > 0) { myers_trace_push(t, sink, 1, t.i); t.i = 0; }
        if (t.j > 0) myers_trace_push(t, sink, 2, t.j);
        myers_trace_push(t, sink, -2, 0);
    }
};

// ---- the walk without a CIGAR: what the coverage counter needs from it, as bit masks over read positions ----
// benchmark_coverage.h:26-67 with threshold 3 keeps the read positions under '=' runs of length >= 3.  The walk
// consumes the read rows in descending order, so it just records, per consumed row p (0-based position),
//   E[p] = the row was consumed by an '=' step,   B[p] = a D step happened between row p and the row above it
// (a D does not consume a row but ends a run), 32 positions per word, accumulated in registers and flushed when a
// word is complete.  A finished word goes to the alignment's 2W slots in shared memory with ONE store (the kernel is
// bound by the ALU pipe, the load/store pipe is nearly idle; a run-time index into a register array would mean local
// memory, a statically unrolled select 4W ALU instructions on almost every step of a warp) and E, B are read back
// with compile-time indices when the walk is over.
// No run lengths, no per-run branches; metrics.cuh:runs3_positions turns E, B into the positions.
struct alignas(8) CoverWord { uint32_t e, b; };
template <int W>
struct CoverTrace {
    int i, j, b;
    uint32_t e, brk;       // word under construction
    uint32_t pend;         // a D step since the last consumed row
    CoverWord* slots;      // word w at slots[w * slot_stride]
    int slot_stride;
    uint32_t E[2 * W], B[2 * W];   // valid after finish()
    GASMA_HD bool live() const { return i > 0 && j > 0; }
    GASMA_HD int col() const { return j; }
    GASMA_HD void start(int m, int n, int hi) {
        i = m; j = n; b = m - n + hi;
        e = brk = pend = 0u;
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
        for (int w = 0; w < 2 * W; ++w) slots[w * slot_stride] = CoverWord{0u, 0u};
    }
    GASMA_HD void land(int w) {   // the word under construction is word w
        slots[w * slot_stride] = CoverWord{e, brk};
        e = brk = 0u;
    }
    GASMA_HD void segment(int j_lo, const TraceCol* tb, int stride) {
        const TraceCol* col = tb + (size_t)(j - 1 - j_lo) * stride;
        while (i > 0 && j > j_lo) {
            const TraceCol c = *col;
            const uint32_t dg = (c.dg >> (b & 31)) & 1u, sel = (c.sel >> (b & 31)) & 1u;
            const uint32_t di = dg | sel, dj = dg | (sel ^ 1u);   // diagonal: both; up (I): row only; left (D): column only
            const int p = i - 1;
            const uint32_t bit = di << (p & 31);
            e |= (dg & sel) ? bit : 0u;
            brk |= pend ? bit : 0u;
            pend = di ^ 1u;
            if (di && (p & 31) == 0) land(p >> 5);
            i -= (int)di;
            j -= (int)dj;
            b += (int)dj - (int)di;
            col -= dj ? stride : 0;
        }
    }
    GASMA_HD void finish() {
        if (i > 0) land((i - 1) >> 5);   // rows left at column 0 are an I run; the word under construction still has to land
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
        for (int w = 0; w < 2 * W; ++w) {
            const CoverWord c = slots[w * slot_stride];
            E[w] = c.e;
            B[w] = c.b;
        }
    }
};

// Penalty and walk of one alignment whose band fits the window.  Only 1/kTraceParts of the columns is resident at a
// time (tb holds myers_trace_slots(n) entries per alignment).  Pass 1 runs the columns once, part by part, leaving a
// checkpoint of the recurrence (8 bytes) at every part boundary and keeping only the last part's windows; that part
// is walked, then each earlier part is recomputed FROM ITS CHECKPOINT into the same slots and walked.  Column work:
// n + n(K-1)/K instead of the n(K+1)/2 of recomputing from column 0, so K can be large: shared memory per alignment
// is 8(ceil(n/K) + K - 1) bytes and that, not the column work, decides how many alignments an SM holds.
// ck: K-1 checkpoints, entry p-1 (state before part p) at ck[(p-1) * ck_stride].  One rolled loop over the 2K-1
// tasks keeps a single instance of the column code in the kernel.
#ifndef GASMA_TRACE_PARTS
#define GASMA_TRACE_PARTS 4   // measured per 1 M pairs (L=100, e=.20, independent pairs): 3 / 4 / 5 / 6 parts = 0.591 / 0.580 / 0.599 / 0.609 ms
#endif
constexpr int kTraceParts = GASMA_TRACE_PARTS;
GASMA_HD int myers_part_start(int n, int part) { return (int)((unsigned)(n * part) / (unsigned)kTraceParts); }   // n <= 256
GASMA_HD constexpr int myers_trace_slots(int max_n) { return (max_n + kTraceParts - 1) / kTraceParts; }
struct alignas(8) BandCk { uint32_t pv, mv; };   // the diagonal-step count is only needed by the first pass (score)
template <int W, class Walker>
GASMA_HD int myers_band_walk(const uint64_t A0[W], const uint64_t A1[W], const uint64_t B0[W], const uint64_t B1[W],
                             int m, int n, int hi, TraceCol* tb, int stride, BandCk* ck, int ck_stride, Walker& wk) {
    constexpr int K = kTraceParts;
    wk.start(m, n, hi);
    if (m == 0 || n == 0) { wk.finish(); return imax(m, n); }
    BandCols c = myers_band_start(hi);
    int pen = 0;
#if defined(__CUDA_ARCH__)
#pragma unroll 1
#endif
    for (int task = 0; task < 2 * K - 1; ++task) {
        const bool first_pass = task < K;
        const int part = first_pass ? task : 2 * K - 2 - task;
        const int lo = myers_part_start(n, part);
        int end;
        if (first_pass) {
            end = myers_part_start(n, part + 1);
            if (part > 0) ck[(size_t)(part - 1) * ck_stride] = BandCk{c.pv, c.mv};
        } else {
            if (!wk.live()) break;
            end = wk.col();   // <= the start of the part walked before: the walk may have gone past its first columns
            if (part > 0) { const BandCk k = ck[(size_t)(part - 1) * ck_stride]; c = BandCols{k.pv, k.mv, 0}; }
            else c = myers_band_start(hi);
        }
        const bool store = task >= K - 1;
        myers_band_run<W>(A0, A1, B0, B1, hi, lo, end, c, store ? tb : nullptr, stride);
        if (task == K - 1) pen = myers_band_score(c, m, n, hi);
        if (store) wk.segment(lo, tb, stride);
    }
    wk.finish();
    return pen;
}

// penalty + CIGAR ops (tb == nullptr: penalty only)
template <int W, class Sink>
GASMA_HD int myers_band_align(const uint64_t A0[W], const uint64_t A1[W], const uint64_t B0[W], const uint64_t B1[W],
                              int m, int n, int hi, TraceCol* tb, int stride, BandCk* ck, int ck_stride, Sink& sink) {
    if (!tb) return (m == 0 || n == 0) ? imax(m, n) : myers_band_forward<W>(A0, A1, B0, B1, m, hi, n);
    OpsWalker<Sink> wk{MyersTrace{}, sink};
    return myers_band_walk<W>(A0, A1, B0, B1, m, n, hi, tb, stride, ck, ck_stride, wk);
}
// penalty + E/B masks
template <int W>
GASMA_HD int myers_band_cover(const uint64_t A0[W], const uint64_t A1[W], const uint64_t B0[W], const uint64_t B1[W],
                              int m, int n, int hi, TraceCol* tb, int stride, BandCk* ck, int ck_stride, CoverWord* slots,
                              int slot_stride, CoverTrace<W>& t) {
    t.slots = slots;
    t.slot_stride = slot_stride;
    return myers_band_walk<W>(A0, A1, B0, B1, m, n, hi, tb, stride, ck, ck_stride, t);
}

}  // namespace gasma
