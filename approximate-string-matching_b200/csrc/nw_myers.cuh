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
// Traceback storage.  The verifier runs in two passes: a penalty-only pass (no memory traffic at all), then, when
// the CIGAR is wanted, a second pass that knows the exact optimum.  An optimal path stays on the diagonals
// d = j - i of nw_band(optimum) (nw.cuh), at most optimum+1 of them, so of each column only the rows
// i in [j-hi, j-lo] can be visited; when the band has <= 32 diagonals one 32-bit WINDOW of DG and of UP per column
// (8 bytes) is all the walk back needs, and a whole alignment's traceback fits in shared memory.  Wider bands go
// to the general Gotoh kernel (same traceback rule).
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

// ---- narrow-band variant: 32-bit windows kept in shared memory -----------------------------------------
// With the exact optimum known (a penalty-only pass), nw_band() has at most penalty+1 diagonals; when that is
// <= 32 the traceback needs one 32-bit window of DG and of UP per column: 8 bytes, small enough to keep a whole
// alignment's traceback in shared memory (tb[(j0*2 + kind) * stride], one column of threads per alignment), so the
// walk back reads shared memory instead of chasing one DRAM sector per step.
template <int NW32>
GASMA_HD uint32_t window32_of(const uint32_t v[NW32], int s) {
    const int idx = s >> 5, sh = s & 31;
    uint32_t x0 = 0u, x1 = 0u;
#pragma unroll
    for (int w = 0; w < NW32; ++w) {
        if (w == idx) x0 = v[w];
        if (w == idx + 1) x1 = v[w];
    }
    return funnel32(x0, x1, sh);
}

// Columns [0, j_end) are computed; the windows of columns j >= j_store are kept at slot (j - j_store).
template <int W>
GASMA_HD int myers_forward_w32(const uint64_t A0[W], const uint64_t A1[W], const uint64_t B0[W], const uint64_t B1[W], int m,
                               int j_store, int j_end, uint32_t* tb, int stride, int band_hi) {
    constexpr int NW32 = 2 * W;
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
    for (int j = 0; j < j_end; ++j) {
        if ((j & 63) == 0) {
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
        score += myers_column<NW32>(eq, pv, mv, dg, m);
        if (j >= j_store) {
            const int s0 = myers_window_start(j, band_hi);
            tb[(size_t)(2 * (j - j_store)) * stride] = window32_of<NW32>(dg, s0);
            tb[(size_t)(2 * (j - j_store) + 1) * stride] = window32_of<NW32>(pv, s0);
        }
    }
    return score;
}

// Resumable walk back: emits ops while j > j_lo (and i > 0), reading the windows stored with base column j_store.
struct MyersTrace {
    int i, j, cnt, cur, run;
};
template <int W>
GASMA_HD void myers_trace_segment(const uint64_t A0[W], const uint64_t A1[W], const uint64_t B0[W], const uint64_t B1[W],
                                  MyersTrace& t, int j_lo, int j_store, const uint32_t* tb, int stride, int band_hi,
                                  uint16_t* row, int cap) {
    auto push = [&](int code, int len) {
        if (code == t.cur) { t.run += len; return; }
        if (t.cur >= 0) { ++t.cnt; if (t.cnt <= cap) row[cap - t.cnt] = (uint16_t)((t.run << 3) | t.cur); }
        t.cur = code; t.run = len;
    };
    while (t.i > 0 && t.j > j_lo) {
        const int b = (t.i - 1) - myers_window_start(t.j - 1, band_hi);   // 0..31 on any optimal path
        const int slot = t.j - 1 - j_store;
        if ((tb[(size_t)(2 * slot) * stride] >> b) & 1u) {
            const int mis = (bit_of<W>(A0, t.i - 1) ^ bit_of<W>(B0, t.j - 1)) | (bit_of<W>(A1, t.i - 1) ^ bit_of<W>(B1, t.j - 1));
            push(mis ? 4 : 3, 1);
            --t.i; --t.j;
        } else if ((tb[(size_t)(2 * slot + 1) * stride] >> b) & 1u) {
            push(1, 1);
            --t.i;
        } else {
            push(2, 1);
            --t.j;
        }
    }
}
// the boundary tail (first row / first column) and the flush of the last run; returns the op count
GASMA_HD int myers_trace_finish(MyersTrace& t, uint16_t* row, int cap) {
    auto push = [&](int code, int len) {
        if (code == t.cur) { t.run += len; return; }
        if (t.cur >= 0) { ++t.cnt; if (t.cnt <= cap) row[cap - t.cnt] = (uint16_t)((t.run << 3) | t.cur); }
        t.cur = code; t.run = len;
    };
    if (t.i > 0) push(1, t.i);
    if (t.j > 0) push(2, t.j);
    push(-2, 0);
    return t.cnt;
}

// Whole traceback of one alignment with HALF the columns resident: the upper half of the columns is stored and
// walked first, then the lower half is recomputed into the same slots (1.5x the column work for half the shared
// memory, i.e. twice the alignments in flight per SM).  tb holds 2 * ((n + 1) / 2) words per alignment.
template <int W>
GASMA_HD int myers_traceback_split(const uint64_t A0[W], const uint64_t A1[W], const uint64_t B0[W], const uint64_t B1[W],
                                   int m, int n, uint32_t* tb, int stride, int band_hi, uint16_t* row, int cap) {
    MyersTrace t{m, n, 0, -1, 0};
    if (m > 0 && n > 0) {
        const int split = n / 2;
        myers_forward_w32<W>(A0, A1, B0, B1, m, split, n, tb, stride, band_hi);
        myers_trace_segment<W>(A0, A1, B0, B1, t, split, split, tb, stride, band_hi, row, cap);
        if (t.i > 0 && t.j > 0) {
            myers_forward_w32<W>(A0, A1, B0, B1, m, 0, t.j, tb, stride, band_hi);
            myers_trace_segment<W>(A0, A1, B0, B1, t, 0, 0, tb, stride, band_hi, row, cap);
        }
    }
    return myers_trace_finish(t, row, cap);
}

}  // namespace gasma
