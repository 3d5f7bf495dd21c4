// nw.cuh -- K4: the NW verifier (optimal global affine-gap penalty + traceback), banded.
//
// Replaces benchmark::_run_nw / _run_nw_sse (GASMA/benchmark/benchmark_utils.h:104-150), i.e. parasail's
// nw_trace with parasail_matrix_create("ACGT",0,-x) (:288): match 0, mismatch x, a gap of length g costs
// o+(g-1)e; penalty = -score (:116,:142).  parasail is an un-vendored submodule of the reference, so the
// arithmetic is restated (Gotoh); the optimum is unique, the traceback tie-breaking is ours and equals
// oracle/gasma_oracle.c:go_nw (diagonal, then I, then D; a gap is closed as early as possible).
//
// Band: with an upper bound C on the optimum (the greedy penalty) every optimal path stays on diagonals d
// (d = j-i) with |d| + |d-(n-m)| <= C / min(o,e), because each gap character costs at least min(o,e).
// Restricting the DP to those diagonals is therefore exact (needs o,e >= 1; otherwise the full matrix is used).
//
// One pair per thread; rows are swept in place over two band arrays H[t], I[t] (t = diagonal index inside
// the band) and a running D; per-cell 4-bit traceback codes go to a scratch matrix.
#pragma once
#include "common.cuh"

namespace gasma {

constexpr int kNwInf = 1 << 28;

struct NwBand { int lo, hi; };  // diagonals lo..hi (inclusive), lo <= min(0,n-m), hi >= max(0,n-m)

GASMA_HD NwBand nw_band(int m, int n, int o, int e, int bound) {
    NwBand b{-m, n};
    const int g = imin(o, e);
    if (bound >= 0 && g >= 1) {
        const int dl = n - m;
        const int B = bound / g;
        const int slack = imax(B - iabs(dl), 0) >> 1;
        b.lo = imax(-m, imin(0, dl) - slack);
        b.hi = imin(n, imax(0, dl) + slack);
    }
    return b;
}

struct NwScratch {
    int32_t* H;      // [t * hstride], t in [0, band]   (one extra sentinel entry)
    int32_t* I;
    int hstride;
    uint32_t* tb;    // [(i-1) * tbw + t/8], 4 bits per cell; may be null (penalty only)
    int tbw;
};

// 64-bit window of a W-word vector starting at bit `pos` (pos may be negative or beyond the vector: zeros)
template <int W>
GASMA_HD uint64_t window64(const uint64_t v[W], int pos) {
    const int idx = pos >> 6, off = pos & 63;  // arithmetic shift: floor division
    uint64_t lo = 0, hi = 0;
#pragma unroll
    for (int j = 0; j < W; ++j) {
        if (j == idx) lo = v[j];
        if (j == idx + 1) hi = v[j];
    }
    return off ? (lo >> off) | (hi << (64 - off)) : lo;
}

template <int W>
GASMA_HD int bit_of(const uint64_t v[W], int pos) {
    uint64_t w = 0;
#pragma unroll
    for (int j = 0; j < W; ++j)
        if (j == (pos >> 6)) w = v[j];
    return (int)((w >> (pos & 63)) & 1ull);
}

// Forward pass.  Returns the optimal penalty H[m][n].
template <int W>
GASMA_HD int nw_forward(int x, int o, int e, const uint64_t A0[W], const uint64_t A1[W], const uint64_t B0[W],
                        const uint64_t B1[W], int m, int n, NwBand bd, const NwScratch& S) {
    const int lo = bd.lo, band = bd.hi - bd.lo + 1;
    int32_t* H = S.H;
    int32_t* I = S.I;
    const int hs = S.hstride;
    // row 0: H[0][j] = D[0][j] = o+(j-1)e
    for (int t = 0; t <= band; ++t) {
        const int j = lo + t;
        H[t * hs] = (t < band && j >= 0 && j <= n) ? (j == 0 ? 0 : o + (j - 1) * e) : kNwInf;
        I[t * hs] = kNwInf;
    }
    for (int i = 1; i <= m; ++i) {
        const int c0 = bit_of<W>(A0, i - 1), c1 = bit_of<W>(A1, i - 1);
        const uint64_t k0 = c0 ? ~0ull : 0ull, k1 = c1 ? ~0ull : 0ull;
        uint64_t eq[W];
#pragma unroll
        for (int j = 0; j < W; ++j) eq[j] = ~(B0[j] ^ k0) & ~(B1[j] ^ k1);  // ref positions equal to read[i-1]
        const int tmin = imax(0, -i - lo), tmax = imin(band - 1, n - i - lo);
        const int base = i + lo - 1;  // ref index (0-based) of cell t is base + t
        int hleft = kNwInf, dleft = kNwInf;
        uint64_t win = 0;
        uint32_t tbword = 0;
        for (int t = tmin; t <= tmax; ++t) {
            const int j = i + lo + t;
            if (t == tmin || (t & 63) == 0) win = window64<W>(eq, base + (t & ~63));
            int hn, in_, dn;
            uint32_t code;
            const int hdiag = H[t * hs];
            const int hup = H[(t + 1) * hs], iup = I[(t + 1) * hs];
            if (j == 0) {  // first column: H[i][0] = I[i][0] = o+(i-1)e
                hn = in_ = o + (i - 1) * e;
                dn = kNwInf;
                code = 1u;
            } else {
                const int io = hup + o, ie = iup + e;
                in_ = imin(io, ie);
                const int dop = hleft + o, dex = dleft + e;
                dn = imin(dop, dex);
                const int match = (int)((win >> (t & 63)) & 1ull);
                const int dg = hdiag + (match ? 0 : x);
                hn = imin(dg, imin(in_, dn));
                code = (hn == dg) ? 0u : (hn == in_) ? 1u : 2u;
                code |= (in_ == io) ? 0u : 4u;   // I extended (opening preferred on ties)
                code |= (dn == dop) ? 0u : 8u;   // D extended
            }
            H[t * hs] = hn;
            I[t * hs] = in_;
            hleft = hn;
            dleft = dn;
            if (S.tb) {
                tbword |= code << (4 * (t & 7));
                if ((t & 7) == 7 || t == tmax) {
                    S.tb[(size_t)(i - 1) * S.tbw + (t >> 3)] = tbword;
                    tbword = 0;
                }
            }
        }
    }
    const int tend = (n - m) - lo;
    return H[tend * hs];
}

// Traceback from (m,n); writes run-length ops (codes = 3 '=', 4 'X', 1 'I', 2 'D') into row[cap-cnt .. cap)
// and returns cnt (ops are produced back to front).
template <int W>
GASMA_HD int nw_traceback(const uint64_t A0[W], const uint64_t A1[W], const uint64_t B0[W], const uint64_t B1[W],
                          int m, int n, NwBand bd, const NwScratch& S, uint16_t* row, int cap) {
    int i = m, j = n, st = 0, cnt = 0;
    int cur = -1, run = 0;
    auto push = [&](int code, int len) {
        if (code == cur) { run += len; return; }
        if (cur >= 0) { ++cnt; if (cnt <= cap) row[cap - cnt] = (uint16_t)((run << 3) | cur); }
        cur = code; run = len;
    };
    while (i > 0 && j > 0) {
        const int t = j - i - bd.lo;
        const uint32_t nib = (S.tb[(size_t)(i - 1) * S.tbw + (t >> 3)] >> (4 * (t & 7))) & 15u;
        if (st == 0) {
            const uint32_t src = nib & 3u;
            if (src == 0u) {
                const int mis = (bit_of<W>(A0, i - 1) ^ bit_of<W>(B0, j - 1)) | (bit_of<W>(A1, i - 1) ^ bit_of<W>(B1, j - 1));
                push(mis ? 4 : 3, 1);
                --i; --j;
            } else st = (int)src;
        } else if (st == 1) {
            push(1, 1);
            if (!(nib & 4u)) st = 0;
            --i;
        } else {
            push(2, 1);
            if (!(nib & 8u)) st = 0;
            --j;
        }
    }
    if (i > 0) push(1, i);
    if (j > 0) push(2, j);
    push(-2, 0);  // flush
    return cnt;
}

}  // namespace gasma
