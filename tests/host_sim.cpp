// tests/host_sim.cpp -- TEST INFRASTRUCTURE.
//
// Compiles the per-pair __host__ __device__ code of the CUDA kernels (csrc/pack.cuh, csrc/greedy.cuh,
// csrc/nw.cuh ...) with g++ and drives it exactly the way the kernels do (tile -> 16-byte chunks ->
// shared bit stream -> per-pair window -> traversal), so that the device logic can be checked against
// the oracle on the CPU-only build container.  It is NOT a fallback: nothing in the product links it,
// and the shipped library fails loudly without a GPU.  Built by tests/conftest.py into tests/_build/.
#include <cstdint>
#include <cstring>
#include <vector>

#include "../approximate-string-matching_b200/csrc/common.cuh"
#include "../approximate-string-matching_b200/csrc/pack.cuh"
#include "../approximate-string-matching_b200/csrc/greedy.cuh"
#include "../approximate-string-matching_b200/csrc/nw.cuh"
#include "../approximate-string-matching_b200/csrc/nw_myers.cuh"
#include "../approximate-string-matching_b200/csrc/leap.cuh"
#include "../approximate-string-matching_b200/csrc/metrics.cuh"
#include "../approximate-string-matching_b200/csrc/generate.cuh"

#include <cmath>

using namespace gasma;

namespace {

template <int W>
constexpr int chunks(int tile) { return tile * 4 * W + 2; }
template <int W>
constexpr int elems(int tile) { return ((chunks<W>(tile) + 2 * (2 * W + 2) + 7) / 8) * 8; }

// host twin of stage_tile (kernels.cuh): same chunking, same partial-tail rule
bool stage(const uint8_t* buf, uint32_t base16, uint32_t end, uint32_t total, uint16_t* s0, uint16_t* s1, int cap) {
    int n = (int)((end - base16 + 15u) >> 4);
    bool ok = n <= cap;
    if (!ok) n = cap;
    for (int c = 0; c < n; ++c) {
        uint32_t a = base16 + 16u * (uint32_t)c, w[4] = {0, 0, 0, 0};
        for (uint32_t i = 0; i < 16u && a + i < total; ++i) w[i >> 2] |= (uint32_t)buf[a + i] << (8u * (i & 3u));
        uint32_t p0, p1;
        pack_chunk(w[0], w[1], w[2], w[3], p0, p1);
        s0[c] = (uint16_t)p0;
        s1[c] = (uint16_t)p1;
    }
    return ok;
}

template <int W, bool CORR = false, int FLIP = 0>
int sim_greedy(const KParams& P, const uint8_t* reads, const uint32_t* roff, const uint8_t* refs, const uint32_t* foff,
               int64_t n, int T, int32_t* cost, uint16_t* ops_stage, int stride, uint32_t* nops, uint32_t* status_out,
               int64_t* steps_out) {
    constexpr int BITS = 64 * W;
    const int nl = CORR ? P.nl : 2 * P.k + 1;
    std::vector<uint32_t> lanes((size_t)nl * 2 * W);
    std::vector<typename LaneState<W>::type> state(nl);
    std::vector<uint16_t> r0(elems<W>(T)), r1(elems<W>(T)), f0(elems<W>(T)), f1(elems<W>(T));
    uint32_t status = 0;
    int64_t steps_total = 0;
    for (int64_t p0 = 0; p0 < n; p0 += T) {
        const int np = (int)std::min<int64_t>(T, n - p0);
        std::fill(r0.begin(), r0.end(), 0); std::fill(r1.begin(), r1.end(), 0);
        std::fill(f0.begin(), f0.end(), 0); std::fill(f1.begin(), f1.end(), 0);
        const uint32_t rbase = roff[p0] & ~15u, fbase = foff[p0] & ~15u;
        bool ok = stage(reads, rbase, roff[p0 + np], roff[n], r0.data(), r1.data(), chunks<W>(T));
        ok &= stage(refs, fbase, foff[p0 + np], foff[n], f0.data(), f1.data(), chunks<W>(T));
        for (int t = 0; t < np; ++t) {
            const int m = (int)(roff[p0 + t + 1] - roff[p0 + t]), nn = (int)(foff[p0 + t + 1] - foff[p0 + t]);
            uint32_t st = 0;
            if (!ok) st |= kStatusTile;
            if (m > BITS || nn > BITS) st |= kStatusLength;
            else if (iabs(nn - m) > kMaxK) st |= kStatusBand;
            else if (CORR && (iabs(nn - m) + P.k > kMaxK || 2 * P.k + 1 + iabs(nn - m) > nl)) st |= kStatusBand;
            int c = INT32_MIN, steps = 0;
            uint32_t my = 0;
            if (!st) {
                uint64_t A0[W], A1[W], B0[W], B1[W];
                const int rbit = (int)(roff[p0 + t] - rbase), fbit = (int)(foff[p0 + t] - fbase);
                extract_plane<W>(reinterpret_cast<const uint32_t*>(r0.data()), rbit, m, A0);
                extract_plane<W>(reinterpret_cast<const uint32_t*>(r1.data()), rbit, m, A1);
                extract_plane<W>(reinterpret_cast<const uint32_t*>(f0.data()), fbit, nn, B0);
                extract_plane<W>(reinterpret_cast<const uint32_t*>(f1.data()), fbit, nn, B1);
                LaneStore<W> S{lanes.data(), state.data(), 1};
                // primary row (kRowOps ops) + overflow row, like the kernels; the harness then leaves the complete list in
                // the pair's staging row whichever of the two holds it
                alignas(16) uint16_t prim[kRowOps];
                uint16_t* row = ops_stage ? ops_stage + (size_t)(p0 + t) * stride : nullptr;
                std::vector<uint16_t> big(stride);
                OpSink sink{prim, row ? row : big.data(), stride, 0u, nullptr, 0};
                sink.begin(0u);
                st = greedy_pair<W, CORR, FLIP>(P, A0, A1, B0, B1, m, nn, S, sink, c, steps);
                sink.finish();
                if (row && sink.n <= kRowKeep) for (int i = 0; i < sink.n; ++i) row[i] = prim[i];
                my = (uint32_t)sink.n;
            }
            cost[p0 + t] = c;
            if (nops) nops[p0 + t] = my;
            status |= st;
            steps_total += steps;
        }
    }
    if (status_out) *status_out = status;
    if (steps_out) *steps_out = steps_total;
    return 0;
}

template <int W>
int sim_pack(const uint8_t* seqs, const uint32_t* off, int64_t n, int T, uint64_t* p0o, uint64_t* p1o, uint32_t* status_out) {
    constexpr int BITS = 64 * W;
    std::vector<uint16_t> s0(elems<W>(T)), s1(elems<W>(T));
    uint32_t status = 0;
    for (int64_t p0 = 0; p0 < n; p0 += T) {
        const int np = (int)std::min<int64_t>(T, n - p0);
        std::fill(s0.begin(), s0.end(), 0); std::fill(s1.begin(), s1.end(), 0);
        const uint32_t base = off[p0] & ~15u;
        bool ok = stage(seqs, base, off[p0 + np], off[n], s0.data(), s1.data(), chunks<W>(T));
        for (int t = 0; t < np; ++t) {
            const int len = (int)(off[p0 + t + 1] - off[p0 + t]);
            uint64_t a[W] = {0}, b[W] = {0};
            if (!ok || len > BITS) status |= kStatusLength;
            else {
                const int bit = (int)(off[p0 + t] - base);
                extract_plane<W>(reinterpret_cast<const uint32_t*>(s0.data()), bit, len, a);
                extract_plane<W>(reinterpret_cast<const uint32_t*>(s1.data()), bit, len, b);
            }
            for (int j = 0; j < W; ++j) { p0o[(p0 + t) * W + j] = a[j]; p1o[(p0 + t) * W + j] = b[j]; }
        }
    }
    if (status_out) *status_out = status;
    return 0;
}

// NW: same staging as the kernels, band arrays with stride 1, traceback rows in a heap scratch
template <int W>
int sim_nw(int x, int o, int e, const uint8_t* reads, const uint32_t* roff, const uint8_t* refs, const uint32_t* foff,
           int64_t n, int T, const int32_t* bound, int32_t* penalty, uint16_t* ops_stage, int stride, uint32_t* nops,
           int cap_band, int64_t* n_wide) {
    constexpr int BITS = 64 * W;
    std::vector<uint16_t> r0(elems<W>(T)), r1(elems<W>(T)), f0(elems<W>(T)), f1(elems<W>(T));
    std::vector<int32_t> H(2 * BITS + 2), I(2 * BITS + 2);
    const int tbw_fast = cap_band / 8, tbw_wide = (2 * BITS + 1 + 7) / 8;
    std::vector<uint32_t> tb((size_t)BITS * tbw_wide);
    int64_t wide = 0;
    for (int64_t p0 = 0; p0 < n; p0 += T) {
        const int np = (int)std::min<int64_t>(T, n - p0);
        std::fill(r0.begin(), r0.end(), 0); std::fill(r1.begin(), r1.end(), 0);
        std::fill(f0.begin(), f0.end(), 0); std::fill(f1.begin(), f1.end(), 0);
        const uint32_t rbase = roff[p0] & ~15u, fbase = foff[p0] & ~15u;
        stage(reads, rbase, roff[p0 + np], roff[n], r0.data(), r1.data(), chunks<W>(T));
        stage(refs, fbase, foff[p0 + np], foff[n], f0.data(), f1.data(), chunks<W>(T));
        for (int t = 0; t < np; ++t) {
            const int64_t p = p0 + t;
            const int m = (int)(roff[p + 1] - roff[p]), nn = (int)(foff[p + 1] - foff[p]);
            uint64_t A0[W], A1[W], B0[W], B1[W];
            extract_plane<W>(reinterpret_cast<const uint32_t*>(r0.data()), (int)(roff[p] - rbase), m, A0);
            extract_plane<W>(reinterpret_cast<const uint32_t*>(r1.data()), (int)(roff[p] - rbase), m, A1);
            extract_plane<W>(reinterpret_cast<const uint32_t*>(f0.data()), (int)(foff[p] - fbase), nn, B0);
            extract_plane<W>(reinterpret_cast<const uint32_t*>(f1.data()), (int)(foff[p] - fbase), nn, B1);
            const NwBand bd = nw_band(m, nn, o, e, bound ? bound[p] : -1);
            const bool is_wide = bd.hi - bd.lo + 1 > cap_band;
            wide += is_wide;
            NwScratch S{H.data(), I.data(), 1, ops_stage ? tb.data() : nullptr, is_wide ? tbw_wide : tbw_fast};
            penalty[p] = nw_forward<W>(x, o, e, A0, A1, B0, B1, m, nn, bd, S);
            if (ops_stage) nops[p] = (uint32_t)nw_traceback<W>(A0, A1, B0, B1, m, nn, bd, S, ops_stage + (size_t)p * stride, stride);
        }
    }
    if (n_wide) *n_wide = wide;
    return 0;
}

// unit-penalty NW the way the kernels run it: the band comes from `bound` (the greedy penalty) or, without one, from
// a penalty-only pass over full columns; a band of <= 32 diagonals takes the one-word banded route (penalty and
// traceback in one pass), anything wider the general Gotoh code
template <int W>
int sim_myers(const uint8_t* reads, const uint32_t* roff, const uint8_t* refs, const uint32_t* foff, int64_t n, int T,
              const int32_t* bound, int32_t* penalty, uint16_t* ops_stage, int stride, uint32_t* nops, int64_t* n_windowed) {
    constexpr int BITS = 64 * W;
    int64_t nwin = 0;
    std::vector<uint16_t> r0(elems<W>(T)), r1(elems<W>(T)), f0(elems<W>(T)), f1(elems<W>(T));
    // exactly the kernel's budget: one part of the columns resident + the checkpoints at the part boundaries
    std::vector<TraceCol> tb((size_t)myers_trace_slots(BITS));
    std::vector<BandCk> ck((size_t)kTraceParts);
    std::vector<uint32_t> tbw((size_t)BITS * ((2 * BITS + 1 + 7) / 8));
    std::vector<int32_t> H(2 * BITS + 2), I(2 * BITS + 2);
    for (int64_t p0 = 0; p0 < n; p0 += T) {
        const int np = (int)std::min<int64_t>(T, n - p0);
        std::fill(r0.begin(), r0.end(), 0); std::fill(r1.begin(), r1.end(), 0);
        std::fill(f0.begin(), f0.end(), 0); std::fill(f1.begin(), f1.end(), 0);
        const uint32_t rbase = roff[p0] & ~15u, fbase = foff[p0] & ~15u;
        stage(reads, rbase, roff[p0 + np], roff[n], r0.data(), r1.data(), chunks<W>(T));
        stage(refs, fbase, foff[p0 + np], foff[n], f0.data(), f1.data(), chunks<W>(T));
        for (int t = 0; t < np; ++t) {
            const int64_t p = p0 + t;
            const int m = (int)(roff[p + 1] - roff[p]), nn = (int)(foff[p + 1] - foff[p]);
            uint64_t A0[W], A1[W], B0[W], B1[W];
            extract_plane<W>(reinterpret_cast<const uint32_t*>(r0.data()), (int)(roff[p] - rbase), m, A0);
            extract_plane<W>(reinterpret_cast<const uint32_t*>(r1.data()), (int)(roff[p] - rbase), m, A1);
            extract_plane<W>(reinterpret_cast<const uint32_t*>(f0.data()), (int)(foff[p] - fbase), nn, B0);
            extract_plane<W>(reinterpret_cast<const uint32_t*>(f1.data()), (int)(foff[p] - fbase), nn, B1);
            int full = -1;
            if (!bound) {
                full = myers_forward<W>(A0, A1, B0, B1, m, nn);
                penalty[p] = full;
                if (!ops_stage) continue;
            }
            NwBand bd = nw_band(m, nn, 1, 1, bound ? bound[p] : full);
            if (!myers_band_fits(bd) && full < 0) {   // loose bound: exact optimum from full columns first
                full = myers_forward<W>(A0, A1, B0, B1, m, nn);
                bd = nw_band(m, nn, 1, 1, full);
            }
            uint16_t* row = ops_stage ? ops_stage + (size_t)p * stride : nullptr;
            if (myers_band_fits(bd)) {
                ++nwin;
                OpsBackSink sink{row ? row + stride : nullptr};
                const int pen = myers_band_align<W>(A0, A1, B0, B1, m, nn, bd.hi, ops_stage ? tb.data() : nullptr, 1, ck.data(), 1, sink);
                if (full >= 0 && pen != full) return -2;
                penalty[p] = pen;
                if (ops_stage) nops[p] = (uint32_t)(row + stride - sink.wp);
            } else {
                NwScratch S{H.data(), I.data(), 1, ops_stage ? tbw.data() : nullptr, (2 * BITS + 1 + 7) / 8};
                const int pen = nw_forward<W>(1, 1, 1, A0, A1, B0, B1, m, nn, bd, S);
                if (full >= 0 && pen != full) return -3;
                penalty[p] = pen;
                if (ops_stage) nops[p] = (uint32_t)nw_traceback<W>(A0, A1, B0, B1, m, nn, bd, S, row, stride);
            }
        }
    }
    if (n_windowed) *n_windowed = nwin;
    return 0;
}

// the benchmark body's fused verifier + counters (nw_band_kernel MODE 2): band from the greedy penalty, the walk's
// '=' runs go straight into the coverage test on 2-bit planes; pairs whose band does not fit take the Gotoh route
// and the character-based coverage test, like nw_wide_kernel + nw_deferred_counters_kernel
template <int W>
int sim_counters(const uint8_t* reads, const uint32_t* roff, const uint8_t* refs, const uint32_t* foff, int64_t n, int T,
                 const int32_t* gcost, const uint16_t* gops, const uint32_t* goff, const int32_t* answers,
                 int32_t* penalty, uint64_t* counters, int64_t* n_deferred) {
    constexpr int BITS = 64 * W;
    const int stride = 2 * BITS + 8;
    std::vector<uint16_t> r0(elems<W>(T)), r1(elems<W>(T)), f0(elems<W>(T)), f1(elems<W>(T)), row(stride);
    std::vector<TraceCol> tb((size_t)myers_trace_slots(BITS));
    std::vector<BandCk> ck((size_t)kTraceParts);
    std::vector<uint32_t> tbw((size_t)BITS * ((2 * BITS + 1 + 7) / 8));
    std::vector<int32_t> H(2 * BITS + 2), I(2 * BITS + 2);
    int64_t ndef = 0;
    for (int64_t p0 = 0; p0 < n; p0 += T) {
        const int np = (int)std::min<int64_t>(T, n - p0);
        std::fill(r0.begin(), r0.end(), 0); std::fill(r1.begin(), r1.end(), 0);
        std::fill(f0.begin(), f0.end(), 0); std::fill(f1.begin(), f1.end(), 0);
        const uint32_t rbase = roff[p0] & ~15u, fbase = foff[p0] & ~15u;
        stage(reads, rbase, roff[p0 + np], roff[n], r0.data(), r1.data(), chunks<W>(T));
        stage(refs, fbase, foff[p0 + np], foff[n], f0.data(), f1.data(), chunks<W>(T));
        for (int t = 0; t < np; ++t) {
            const int64_t p = p0 + t;
            const int m = (int)(roff[p + 1] - roff[p]), nn = (int)(foff[p + 1] - foff[p]);
            uint64_t A0[W], A1[W], B0[W], B1[W];
            extract_plane<W>(reinterpret_cast<const uint32_t*>(r0.data()), (int)(roff[p] - rbase), m, A0);
            extract_plane<W>(reinterpret_cast<const uint32_t*>(r1.data()), (int)(roff[p] - rbase), m, A1);
            extract_plane<W>(reinterpret_cast<const uint32_t*>(f0.data()), (int)(foff[p] - fbase), nn, B0);
            extract_plane<W>(reinterpret_cast<const uint32_t*>(f1.data()), (int)(foff[p] - fbase), nn, B1);
            NwBand bd = nw_band(m, nn, 1, 1, gcost[p]);
            if (!myers_band_fits(bd)) bd = nw_band(m, nn, 1, 1, myers_forward<W>(A0, A1, B0, B1, m, nn));
            int pen, cov;
            if (myers_band_fits(bd)) {
                CoverTrace<W> ct;
                CoverWord eb[2 * W];
                pen = myers_band_cover<W>(A0, A1, B0, B1, m, nn, bd.hi, tb.data(), 1, ck.data(), 1, eb, 1, ct);
                uint32_t P1[2 * W], P2[2 * W], a0[2 * W], a1[2 * W];
                runs3_positions<2 * W>(ct.E, ct.B, P2);
                lcm_mask<2 * W>(gops + goff[p], (int)(goff[p + 1] - goff[p]), 1, P1);   // LCM1: read positions under greedy M runs
                for (int w = 0; w < W; ++w) {
                    a0[2 * w] = (uint32_t)A0[w]; a0[2 * w + 1] = (uint32_t)(A0[w] >> 32);
                    a1[2 * w] = (uint32_t)A1[w]; a1[2 * w + 1] = (uint32_t)(A1[w] >> 32);
                }
                uint32_t outside = 0;
                for (int w = 0; w < 2 * W; ++w) outside |= P2[w] & ~P1[w];
                cov = outside == 0 ? 1 : covers_planes<2 * W>(a0, a1, m, P1, P2);   // kernel: queued for cover_hard_kernel
            } else {
                ++ndef;
                NwScratch S{H.data(), I.data(), 1, tbw.data(), (2 * BITS + 1 + 7) / 8};
                pen = nw_forward<W>(1, 1, 1, A0, A1, B0, B1, m, nn, bd, S);
                const int cnt = nw_traceback<W>(A0, A1, B0, B1, m, nn, bd, S, row.data(), stride);
                uint32_t scratch[2 * kCovWords];
                cov = coverage_pair(reads + roff[p], m, gops + goff[p], (int)(goff[p + 1] - goff[p]), row.data() + stride - cnt, cnt, scratch, 1);
            }
            penalty[p] = pen;
            const int correct = (answers && answers[p] != INT32_MIN) ? answers[p] : pen;
            counters[0] += 1;
            counters[1] += pen == correct;
            counters[2] += gcost[p] == correct;
            counters[3] += (uint64_t)cov;
        }
    }
    if (n_deferred) *n_deferred = ndef;
    return 0;
}

// LEAP comparator the way leap_kernel runs it: staged streams -> planes -> leap_pair on per-pair tables
template <int W>
int sim_leap(int k, int x, int o, int e, const uint8_t* reads, const uint32_t* roff, const uint8_t* refs, const uint32_t* foff,
             int64_t n, int T, int32_t* penalty) {
    const int depth = leap_depth(x, o, e);
    const int nl = (2 * k + 1) * 2 * W, nr = (2 * k + 3) * depth;
    std::vector<uint32_t> words((size_t)leap_words_per_pair(k, 2 * W, depth));
    std::vector<uint16_t> r0(elems<W>(T)), r1(elems<W>(T)), f0(elems<W>(T)), f1(elems<W>(T));
    for (int64_t p0 = 0; p0 < n; p0 += T) {
        const int np = (int)std::min<int64_t>(T, n - p0);
        std::fill(r0.begin(), r0.end(), 0); std::fill(r1.begin(), r1.end(), 0);
        std::fill(f0.begin(), f0.end(), 0); std::fill(f1.begin(), f1.end(), 0);
        const uint32_t rbase = roff[p0] & ~15u, fbase = foff[p0] & ~15u;
        stage(reads, rbase, roff[p0 + np], roff[n], r0.data(), r1.data(), chunks<W>(T));
        stage(refs, fbase, foff[p0 + np], foff[n], f0.data(), f1.data(), chunks<W>(T));
        for (int t = 0; t < np; ++t) {
            const int64_t p = p0 + t;
            const int m = (int)(roff[p + 1] - roff[p]), nn = (int)(foff[p + 1] - foff[p]);
            uint64_t A0[W], A1[W], B0[W], B1[W];
            extract_plane<W>(reinterpret_cast<const uint32_t*>(r0.data()), (int)(roff[p] - rbase), m, A0);
            extract_plane<W>(reinterpret_cast<const uint32_t*>(r1.data()), (int)(roff[p] - rbase), m, A1);
            extract_plane<W>(reinterpret_cast<const uint32_t*>(f0.data()), (int)(foff[p] - fbase), nn, B0);
            extract_plane<W>(reinterpret_cast<const uint32_t*>(f1.data()), (int)(foff[p] - fbase), nn, B1);
            std::fill(words.begin(), words.end(), 0xDEADBEEFu);   // the tables start as garbage on the device too
            LeapStore S{words.data(), reinterpret_cast<int*>(words.data() + nl), reinterpret_cast<int*>(words.data() + nl + nr),
                        reinterpret_cast<int*>(words.data() + nl + 2 * nr), 1, depth};
            penalty[p] = leap_pair<W>(k, x, o, e, A0, A1, B0, B1, m, nn, S);
        }
    }
    return 0;
}

}  // namespace

extern "C" {

int sim_myers_batch(int W, int T, const uint8_t* reads, const uint32_t* roff, const uint8_t* refs, const uint32_t* foff,
                    int64_t n, const int32_t* bound, int32_t* penalty, uint16_t* ops_stage, int stride, uint32_t* nops, int64_t* n_windowed) {
    switch (W) {
        case 1: return sim_myers<1>(reads, roff, refs, foff, n, T, bound, penalty, ops_stage, stride, nops, n_windowed);
        case 2: return sim_myers<2>(reads, roff, refs, foff, n, T, bound, penalty, ops_stage, stride, nops, n_windowed);
        case 3: return sim_myers<3>(reads, roff, refs, foff, n, T, bound, penalty, ops_stage, stride, nops, n_windowed);
        case 4: return sim_myers<4>(reads, roff, refs, foff, n, T, bound, penalty, ops_stage, stride, nops, n_windowed);
    }
    return -1;
}

int sim_counters_batch(int W, int T, const uint8_t* reads, const uint32_t* roff, const uint8_t* refs, const uint32_t* foff,
                       int64_t n, const int32_t* gcost, const uint16_t* gops, const uint32_t* goff, const int32_t* answers,
                       int32_t* penalty, uint64_t* counters, int64_t* n_deferred) {
    switch (W) {
        case 1: return sim_counters<1>(reads, roff, refs, foff, n, T, gcost, gops, goff, answers, penalty, counters, n_deferred);
        case 2: return sim_counters<2>(reads, roff, refs, foff, n, T, gcost, gops, goff, answers, penalty, counters, n_deferred);
        case 3: return sim_counters<3>(reads, roff, refs, foff, n, T, gcost, gops, goff, answers, penalty, counters, n_deferred);
        case 4: return sim_counters<4>(reads, roff, refs, foff, n, T, gcost, gops, goff, answers, penalty, counters, n_deferred);
    }
    return -1;
}

int sim_leap_batch(int W, int T, int k, int x, int o, int e, const uint8_t* reads, const uint32_t* roff, const uint8_t* refs,
                   const uint32_t* foff, int64_t n, int32_t* penalty) {
    switch (W) {
        case 1: return sim_leap<1>(k, x, o, e, reads, roff, refs, foff, n, T, penalty);
        case 2: return sim_leap<2>(k, x, o, e, reads, roff, refs, foff, n, T, penalty);
        case 3: return sim_leap<3>(k, x, o, e, reads, roff, refs, foff, n, T, penalty);
        case 4: return sim_leap<4>(k, x, o, e, reads, roff, refs, foff, n, T, penalty);
    }
    return -1;
}

int sim_nw_batch(int W, int T, int x, int o, int e, const uint8_t* reads, const uint32_t* roff, const uint8_t* refs,
                 const uint32_t* foff, int64_t n, const int32_t* bound, int32_t* penalty, uint16_t* ops_stage, int stride,
                 uint32_t* nops, int cap_band, int64_t* n_wide) {
    switch (W) {
        case 1: return sim_nw<1>(x, o, e, reads, roff, refs, foff, n, T, bound, penalty, ops_stage, stride, nops, cap_band, n_wide);
        case 2: return sim_nw<2>(x, o, e, reads, roff, refs, foff, n, T, bound, penalty, ops_stage, stride, nops, cap_band, n_wide);
        case 3: return sim_nw<3>(x, o, e, reads, roff, refs, foff, n, T, bound, penalty, ops_stage, stride, nops, cap_band, n_wide);
        case 4: return sim_nw<4>(x, o, e, reads, roff, refs, foff, n, T, bound, penalty, ops_stage, stride, nops, cap_band, n_wide);
    }
    return -1;
}

int sim_coverage(const uint8_t* s1, int s1_len, const uint16_t* g_ops, int g_n, const uint16_t* n_ops, int n_n) {
    uint32_t scratch[2 * kCovWords];
    return coverage_pair(s1, s1_len, g_ops, g_n, n_ops, n_n, scratch, 1);
}

int sim_generate_pair(uint64_t seed, int length, int mode, int num_errors, uint64_t thr_mismatch, uint64_t thr_sub,
                      uint64_t thr_ins, uint64_t thr_del, uint64_t pair, char* read, char* ref, int* len_only) {
    GenCfg c{seed, length, mode, num_errors, thr_mismatch, thr_sub, thr_ins, thr_del};
    *len_only = gen_pair<false>(c, pair, nullptr, nullptr);
    return gen_pair<true>(c, pair, read, ref);
}

int sim_greedy_batch(int W, int T, int k, int x, int o, int e, int global, double pm, double pmm, double pid,
                     const uint8_t* reads, const uint32_t* roff, const uint8_t* refs, const uint32_t* foff, int64_t n,
                     int32_t* cost, uint16_t* ops_stage, int stride, uint32_t* nops, uint32_t* status, int64_t* steps) {
    KParams P{k, x, o, e, global, std::log(pm / 0.25), std::log(pmm / 0.25), std::log(pid / 2 / 0.25)};
    switch (W) {
        case 1: return sim_greedy<1>(P, reads, roff, refs, foff, n, T, cost, ops_stage, stride, nops, status, steps);
        case 2: return sim_greedy<2>(P, reads, roff, refs, foff, n, T, cost, ops_stage, stride, nops, status, steps);
        case 3: return sim_greedy<3>(P, reads, roff, refs, foff, n, T, cost, ops_stage, stride, nops, status, steps);
        case 4: return sim_greedy<4>(P, reads, roff, refs, foff, n, T, cost, ops_stage, stride, nops, status, steps);
    }
    return -1;
}

// the widened band of the reference's -DCORRECTION build (gasma_set_correction); nl = 2k+1+max_delta lane slots
int sim_greedy_batch_corr(int W, int T, int k, int x, int o, int e, int global, double pm, double pmm, double pid, int max_delta,
                          const uint8_t* reads, const uint32_t* roff, const uint8_t* refs, const uint32_t* foff, int64_t n,
                          int32_t* cost, uint16_t* ops_stage, int stride, uint32_t* nops, uint32_t* status, int64_t* steps) {
    KParams P{k, x, o, e, global, std::log(pm / 0.25), std::log(pmm / 0.25), std::log(pid / 2 / 0.25), 1, 2 * k + 1 + max_delta};
    switch (W) {
        case 1: return sim_greedy<1, true>(P, reads, roff, refs, foff, n, T, cost, ops_stage, stride, nops, status, steps);
        case 2: return sim_greedy<2, true>(P, reads, roff, refs, foff, n, T, cost, ops_stage, stride, nops, status, steps);
        case 3: return sim_greedy<3, true>(P, reads, roff, refs, foff, n, T, cost, ops_stage, stride, nops, status, steps);
        case 4: return sim_greedy<4, true>(P, reads, roff, refs, foff, n, T, cost, ops_stage, stride, nops, status, steps);
    }
    return -1;
}

// the aligner's other mask variants (gasma_set_mask_variant): 1 = flip_short_hurdles(2), 2 = + flip_short_matches(1)
int sim_greedy_batch_mask(int variant, int W, int T, int k, int x, int o, int e, int global, double pm, double pmm, double pid,
                          const uint8_t* reads, const uint32_t* roff, const uint8_t* refs, const uint32_t* foff, int64_t n,
                          int32_t* cost, uint16_t* ops_stage, int stride, uint32_t* nops, uint32_t* status, int64_t* steps) {
    KParams P{k, x, o, e, global, std::log(pm / 0.25), std::log(pmm / 0.25), std::log(pid / 2 / 0.25)};
#define GASMA_SIM_CASE(WW, FF) if (W == WW && variant == FF) return sim_greedy<WW, false, FF>(P, reads, roff, refs, foff, n, T, cost, ops_stage, stride, nops, status, steps)
    GASMA_SIM_CASE(1, 1); GASMA_SIM_CASE(2, 1); GASMA_SIM_CASE(3, 1); GASMA_SIM_CASE(4, 1);
    GASMA_SIM_CASE(1, 2); GASMA_SIM_CASE(2, 2); GASMA_SIM_CASE(3, 2); GASMA_SIM_CASE(4, 2);
#undef GASMA_SIM_CASE
    return -1;
}

int sim_pack_batch(int W, int T, const uint8_t* seqs, const uint32_t* off, int64_t n, uint64_t* p0, uint64_t* p1,
                   uint32_t* status) {
    switch (W) {
        case 1: return sim_pack<1>(seqs, off, n, T, p0, p1, status);
        case 2: return sim_pack<2>(seqs, off, n, T, p0, p1, status);
        case 3: return sim_pack<3>(seqs, off, n, T, p0, p1, status);
        case 4: return sim_pack<4>(seqs, off, n, T, p0, p1, status);
    }
    return -1;
}

}  // extern "C"
