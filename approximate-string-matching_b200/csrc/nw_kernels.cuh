// nw_kernels.cuh -- __global__ entry points for K4 (NW verifier) and K5 (accuracy / coverage counters).
#pragma once
#include "kernels.cuh"
#include "nw.cuh"
#include "nw_myers.cuh"
#include "metrics.cuh"

namespace gasma {

constexpr int kNwCap = 64;   // widest band (diagonals) the shared-memory kernel handles; wider pairs are deferred

template <int W>
__host__ __device__ constexpr size_t nw_smem_bytes(int tile) {
    return 2 * (size_t)(kNwCap + 1) * tile * 4            // H, I band arrays
           + 4 * (size_t)stream_elems<W>(tile) * 2        // plane streams
           + 2 * (size_t)(tile + 1) * 4 + 64;
}
template <int W>
__host__ __device__ constexpr int nw_tb_words_fast() { return kNwCap / 8; }
template <int W>
__host__ __device__ constexpr int nw_tb_words_wide() { return (2 * 64 * W + 1 + 7) / 8; }

// Fast kernel: one CTA per tile, one thread per pair, band arrays in shared memory.
// tb: per-pair traceback scratch, pair p of this launch uses tb + (p - first) * 64W * tbw words.
template <int W, int T>
__global__ void __launch_bounds__(T)
nw_kernel(int x, int o, int e, const uint8_t* __restrict__ reads, const uint32_t* __restrict__ read_off,
          const uint8_t* __restrict__ refs, const uint32_t* __restrict__ ref_off, long long n,
          long long first, long long count,   // this launch covers pairs [first, first+count)
          const int32_t* __restrict__ bound, int32_t* __restrict__ penalty, uint32_t* __restrict__ tb,
          uint16_t* __restrict__ ops_stage, int ops_stride, uint32_t* __restrict__ nops_out,
          uint32_t* __restrict__ deferred, uint32_t* __restrict__ deferred_count, uint32_t* __restrict__ status) {
    constexpr int BITS = 64 * W;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    int32_t* s_H = reinterpret_cast<int32_t*>(smem_raw);
    int32_t* s_I = s_H + (kNwCap + 1) * T;
    uint16_t* s_r0 = reinterpret_cast<uint16_t*>(s_I + (kNwCap + 1) * T);
    uint16_t* s_r1 = s_r0 + stream_elems<W>(T);
    uint16_t* s_f0 = s_r1 + stream_elems<W>(T);
    uint16_t* s_f1 = s_f0 + stream_elems<W>(T);
    uint32_t* s_roff = reinterpret_cast<uint32_t*>(s_f1 + stream_elems<W>(T));
    uint32_t* s_foff = s_roff + (T + 1);

    const int tid = threadIdx.x;
    const long long p0 = first + (long long)blockIdx.x * T;
    const long long pend = first + count;
    const int np = (int)((pend - p0) < (long long)T ? (pend - p0) : (long long)T);
    for (int i = tid; i <= np; i += T) { s_roff[i] = read_off[p0 + i]; s_foff[i] = ref_off[p0 + i]; }
    for (int i = tid; i < stream_elems<W>(T); i += T) { s_r0[i] = 0; s_r1[i] = 0; s_f0[i] = 0; s_f1[i] = 0; }
    __syncthreads();
    const uint32_t rbase = s_roff[0] & ~15u, fbase = s_foff[0] & ~15u;
    bool ok = stage_tile(reads, rbase, s_roff[np], read_off[n], s_r0, s_r1, stream_chunks<W>(T), tid, T);
    ok &= stage_tile(refs, fbase, s_foff[np], ref_off[n], s_f0, s_f1, stream_chunks<W>(T), tid, T);
    __syncthreads();

    uint32_t st = 0;
    if (tid < np) {
        const long long p = p0 + tid;
        const int m = (int)(s_roff[tid + 1] - s_roff[tid]);
        const int nn = (int)(s_foff[tid + 1] - s_foff[tid]);
        if (!ok) st |= kStatusTile;
        if (m > BITS || nn > BITS) st |= kStatusLength;
        int pen = INT32_MIN;
        uint32_t cnt = 0;
        if (st == 0) {
            const NwBand bd = nw_band(m, nn, o, e, bound ? bound[p] : -1);
            if (bd.hi - bd.lo + 1 > kNwCap) {
                deferred[atomicAdd(deferred_count, 1u)] = (uint32_t)p;   // handled by nw_wide_kernel
            } else {
                uint64_t A0[W], A1[W], B0[W], B1[W];
                const int rbit = (int)(s_roff[tid] - rbase), fbit = (int)(s_foff[tid] - fbase);
                extract_plane<W>(reinterpret_cast<const uint32_t*>(s_r0), rbit, m, A0);
                extract_plane<W>(reinterpret_cast<const uint32_t*>(s_r1), rbit, m, A1);
                extract_plane<W>(reinterpret_cast<const uint32_t*>(s_f0), fbit, nn, B0);
                extract_plane<W>(reinterpret_cast<const uint32_t*>(s_f1), fbit, nn, B1);
                NwScratch S{s_H + tid, s_I + tid, T,
                            ops_stage ? tb + (size_t)(p - first) * BITS * nw_tb_words_fast<W>() : nullptr, nw_tb_words_fast<W>()};
                pen = nw_forward<W>(x, o, e, A0, A1, B0, B1, m, nn, bd, S);
                if (ops_stage)
                    cnt = (uint32_t)nw_traceback<W>(A0, A1, B0, B1, m, nn, bd, S, ops_stage + (size_t)p * ops_stride, ops_stride);
            }
        }
        penalty[p] = pen;
        if (nops_out) nops_out[p] = cnt;
    }
    const uint32_t wst = __reduce_or_sync(0xffffffffu, st);
    if (wst && (tid & 31) == 0) atomicOr(status, wst);
}

// Wide kernel: pairs whose band exceeds kNwCap (or all pairs when no bound is given).  Band arrays and
// traceback live in a pool of global-memory slots; each thread owns one slot and walks the deferred list.
template <int W>
__global__ void __launch_bounds__(128)
nw_wide_kernel(int x, int o, int e, const uint8_t* __restrict__ reads, const uint32_t* __restrict__ read_off,
               const uint8_t* __restrict__ refs, const uint32_t* __restrict__ ref_off,
               const int32_t* bound, int32_t* penalty,   // may alias (pass 2 of the unit-penalty verifier)
               int32_t* __restrict__ pool_hi, uint32_t* __restrict__ pool_tb, int nslots,
               uint16_t* __restrict__ ops_stage, int ops_stride, uint32_t* __restrict__ nops_out,
               const uint32_t* __restrict__ deferred, const uint32_t* __restrict__ deferred_count) {
    constexpr int BITS = 64 * W;
    constexpr int HN = 2 * BITS + 2;
    const int slot = blockIdx.x * blockDim.x + threadIdx.x;
    if (slot >= nslots) return;
    const uint32_t total = *deferred_count;
    for (uint32_t it = (uint32_t)slot; it < total; it += (uint32_t)nslots) {
        const uint32_t p = deferred[it];
        const uint32_t r0 = read_off[p], f0 = ref_off[p];
        const int m = (int)(read_off[p + 1] - r0), nn = (int)(ref_off[p + 1] - f0);
        // pack this pair's planes straight from global memory (scalar path: the wide kernel is the slow path)
        uint64_t A0[W], A1[W], B0[W], B1[W];
#pragma unroll
        for (int j = 0; j < W; ++j) A0[j] = A1[j] = B0[j] = B1[j] = 0ull;
        for (int i = 0; i < m; ++i) {
            const uint8_t c = reads[r0 + i];
            const uint64_t b0 = (c == 'C' || c == 'T'), b1 = (c == 'G' || c == 'T');
#pragma unroll
            for (int j = 0; j < W; ++j) if (j == (i >> 6)) { A0[j] |= b0 << (i & 63); A1[j] |= b1 << (i & 63); }
        }
        for (int i = 0; i < nn; ++i) {
            const uint8_t c = refs[f0 + i];
            const uint64_t b0 = (c == 'C' || c == 'T'), b1 = (c == 'G' || c == 'T');
#pragma unroll
            for (int j = 0; j < W; ++j) if (j == (i >> 6)) { B0[j] |= b0 << (i & 63); B1[j] |= b1 << (i & 63); }
        }
        const NwBand bd = nw_band(m, nn, o, e, bound ? bound[p] : -1);
        NwScratch S{pool_hi + (size_t)slot * 2 * HN, pool_hi + (size_t)slot * 2 * HN + HN, 1,
                    ops_stage ? pool_tb + (size_t)slot * BITS * nw_tb_words_wide<W>() : nullptr, nw_tb_words_wide<W>()};
        penalty[p] = nw_forward<W>(x, o, e, A0, A1, B0, B1, m, nn, bd, S);
        if (ops_stage) {
            const int cnt = nw_traceback<W>(A0, A1, B0, B1, m, nn, bd, S, ops_stage + (size_t)p * ops_stride, ops_stride);
            if (nops_out) nops_out[p] = (uint32_t)cnt;
        }
    }
}

// Unit-penalty verifier (x = o = e = 1), pass 1: Myers bit-vector columns, penalty only, one pair per thread.
template <int W>
__host__ __device__ constexpr size_t myers_smem_bytes(int tile) {
    return 4 * (size_t)stream_elems<W>(tile) * 2 + 2 * (size_t)(tile + 1) * 4 + 64;
}
// the banded kernel adds the per-alignment traceback windows: 1/kTraceParts of the 64W columns at a time x 8 bytes
template <int W>
__host__ __device__ constexpr size_t myers_trace_smem_bytes(int tile) {
    return myers_smem_bytes<W>(tile) + (size_t)myers_trace_slots(64 * W) * sizeof(TraceCol) * tile;
}

template <int W, int T>
__global__ void __launch_bounds__(T)
nw_myers_kernel(const uint8_t* __restrict__ reads, const uint32_t* __restrict__ read_off, const uint8_t* __restrict__ refs,
                const uint32_t* __restrict__ ref_off, long long n, int32_t* __restrict__ penalty, uint32_t* __restrict__ status) {
    constexpr int BITS = 64 * W;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    uint16_t* s_r0 = reinterpret_cast<uint16_t*>(smem_raw);
    uint16_t* s_r1 = s_r0 + stream_elems<W>(T);
    uint16_t* s_f0 = s_r1 + stream_elems<W>(T);
    uint16_t* s_f1 = s_f0 + stream_elems<W>(T);
    uint32_t* s_roff = reinterpret_cast<uint32_t*>(s_f1 + stream_elems<W>(T));
    uint32_t* s_foff = s_roff + (T + 1);

    const int tid = threadIdx.x;
    const long long p0 = (long long)blockIdx.x * T;
    const int np = (int)((n - p0) < (long long)T ? (n - p0) : (long long)T);
    for (int i = tid; i <= np; i += T) { s_roff[i] = read_off[p0 + i]; s_foff[i] = ref_off[p0 + i]; }
    for (int i = tid; i < stream_elems<W>(T); i += T) { s_r0[i] = 0; s_r1[i] = 0; s_f0[i] = 0; s_f1[i] = 0; }
    __syncthreads();
    const uint32_t rbase = s_roff[0] & ~15u, fbase = s_foff[0] & ~15u;
    bool ok = stage_tile(reads, rbase, s_roff[np], read_off[n], s_r0, s_r1, stream_chunks<W>(T), tid, T);
    ok &= stage_tile(refs, fbase, s_foff[np], ref_off[n], s_f0, s_f1, stream_chunks<W>(T), tid, T);
    __syncthreads();

    uint32_t st = 0;
    if (tid < np) {
        const long long p = p0 + tid;
        const int m = (int)(s_roff[tid + 1] - s_roff[tid]);
        const int nn = (int)(s_foff[tid + 1] - s_foff[tid]);
        if (!ok) st |= kStatusTile;
        if (m > BITS || nn > BITS) st |= kStatusLength;
        int pen = INT32_MIN;
        if (st == 0) {
            uint64_t A0[W], A1[W], B0[W], B1[W];
            const int rbit = (int)(s_roff[tid] - rbase), fbit = (int)(s_foff[tid] - fbase);
            extract_plane<W>(reinterpret_cast<const uint32_t*>(s_r0), rbit, m, A0);
            extract_plane<W>(reinterpret_cast<const uint32_t*>(s_r1), rbit, m, A1);
            extract_plane<W>(reinterpret_cast<const uint32_t*>(s_f0), fbit, nn, B0);
            extract_plane<W>(reinterpret_cast<const uint32_t*>(s_f1), fbit, nn, B1);
            pen = myers_forward<W>(A0, A1, B0, B1, m, nn);
        }
        penalty[p] = pen;
    }
    const uint32_t wst = __reduce_or_sync(0xffffffffu, st);
    if (wst && (tid & 31) == 0) atomicOr(status, wst);
}

// Banded verifier: one 32-bit window per column (nw_myers.cuh), penalty and traceback in one pass.
// `bound` is an upper bound on every pair's optimum: the greedy penalty, or the exact optimum from nw_myers_kernel
// (it may then alias `penalty`).  Pairs whose band has more than 32 diagonals are appended to `deferred` for
// nw_wide_kernel (general Gotoh, same traceback rule) with their exact optimum in `penalty` as its bound.
//   MODE 0: penalty only (no shared-memory windows)
//   MODE 1: penalty + CIGAR ops into the pair's staging row (back to front) + nops_out
//   MODE 2: penalty + the benchmark's counters (metrics.cuh): the walk hands its '=' steps to the coverage test
//           as bit masks, no NW CIGAR is written.  bound = the greedy penalty, g_ops/g_off = the greedy CIGAR CSR.
//           A deferred pair is not counted here: nw_wide_kernel writes its CIGAR and
//           nw_deferred_counters_kernel counts it from there.
constexpr int kNwPenalty = 0, kNwOps = 1, kNwCounters = 2;
// MODE 2 keeps the greedy CIGAR ops of the tile's pairs (one contiguous CSR range) in shared memory, copied with
// coalesced loads next to the tile staging: the per-pair loop over them then runs on shared-memory latency instead
// of one dependent L2 round trip per op at a handful of active threads.  A tile with more ops reads them from global.
constexpr int kTileGops = 512;

struct NwCounterArgs {
    const uint16_t* g_ops;        // greedy CIGAR (CSR)
    const uint32_t* g_off;
    const int32_t* answers;       // optional (read_answer_file); INT32_MIN or null = use the NW penalty
    unsigned long long* counters; // total, nw_correct, greedy_correct, greedy_coverage
    uint32_t* hard;               // [8W+1][hard_cap]: pairs whose coverage needs the sequential test (cover_hard_kernel)
    uint32_t* hard_count;
    uint32_t hard_cap;
};

template <int W, int T, int MODE>
__global__ void __launch_bounds__(T)
nw_band_kernel(const uint8_t* __restrict__ reads, const uint32_t* __restrict__ read_off, const uint8_t* __restrict__ refs,
               const uint32_t* __restrict__ ref_off, long long n, const int32_t* bound, int32_t* penalty,
               uint16_t* __restrict__ ops_stage, int ops_stride, uint32_t* __restrict__ nops_out,
               uint32_t* __restrict__ deferred, uint32_t* __restrict__ deferred_count, uint32_t* __restrict__ status,
               NwCounterArgs ca) {
    constexpr int BITS = 64 * W;
    constexpr bool TRACE = MODE != kNwPenalty;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    TraceCol* s_tb = reinterpret_cast<TraceCol*>(smem_raw);                       // [col][T]
    uint16_t* s_r0 = reinterpret_cast<uint16_t*>(s_tb + (TRACE ? (size_t)myers_trace_slots(BITS) * T : 0));
    uint16_t* s_r1 = s_r0 + stream_elems<W>(T);
    uint16_t* s_f0 = s_r1 + stream_elems<W>(T);
    uint16_t* s_f1 = s_f0 + stream_elems<W>(T);
    uint32_t* s_roff = reinterpret_cast<uint32_t*>(s_f1 + stream_elems<W>(T));
    uint32_t* s_foff = s_roff + (T + 1);
    uint16_t* s_gops = reinterpret_cast<uint16_t*>(s_foff + (T + 1));   // MODE 2 only (kTileGops entries)

    const int tid = threadIdx.x;
    const long long p0 = (long long)blockIdx.x * T;
    const int np = (int)((n - p0) < (long long)T ? (n - p0) : (long long)T);
    uint32_t g_lo = 0u;
    bool gops_staged = false;
    if (MODE == kNwCounters) {
        g_lo = ca.g_off[p0];
        const uint32_t g_cnt = ca.g_off[p0 + np] - g_lo;
        gops_staged = g_cnt <= (uint32_t)kTileGops;
        if (gops_staged)
            for (uint32_t i = tid; i < g_cnt; i += T) s_gops[i] = ca.g_ops[g_lo + i];
    }
    for (int i = tid; i <= np; i += T) { s_roff[i] = read_off[p0 + i]; s_foff[i] = ref_off[p0 + i]; }
    for (int i = tid; i < stream_elems<W>(T); i += T) { s_r0[i] = 0; s_r1[i] = 0; s_f0[i] = 0; s_f1[i] = 0; }
    __syncthreads();
    const uint32_t rbase = s_roff[0] & ~15u, fbase = s_foff[0] & ~15u;
    bool ok = stage_tile(reads, rbase, s_roff[np], read_off[n], s_r0, s_r1, stream_chunks<W>(T), tid, T);
    ok &= stage_tile(refs, fbase, s_foff[np], ref_off[n], s_f0, s_f1, stream_chunks<W>(T), tid, T);
    __syncthreads();

    uint32_t st = 0;
    uint32_t c_tot = 0, c_nw = 0, c_gr = 0, c_cov = 0;   // MODE 2
    if (tid < np) {
        const long long p = p0 + tid;
        const int m = (int)(s_roff[tid + 1] - s_roff[tid]);
        const int nn = (int)(s_foff[tid + 1] - s_foff[tid]);
        if (!ok) st |= kStatusTile;
        if (m > BITS || nn > BITS) st |= kStatusLength;
        const int bnd = bound[p];
        int cnt = 0;
        if (st != 0 || bnd == INT32_MIN) {   // INT32_MIN: rejected by the pass that produced the bound
            penalty[p] = INT32_MIN;
        } else {
            uint64_t A0[W], A1[W], B0[W], B1[W];
            const int rbit = (int)(s_roff[tid] - rbase), fbit = (int)(s_foff[tid] - fbase);
            extract_plane<W>(reinterpret_cast<const uint32_t*>(s_r0), rbit, m, A0);
            extract_plane<W>(reinterpret_cast<const uint32_t*>(s_r1), rbit, m, A1);
            extract_plane<W>(reinterpret_cast<const uint32_t*>(s_f0), fbit, nn, B0);
            extract_plane<W>(reinterpret_cast<const uint32_t*>(s_f1), fbit, nn, B1);
            NwBand bd = nw_band(m, nn, 1, 1, bnd);
            if (!myers_band_fits(bd)) {
                // a loose bound (rare: the greedy penalty is usually close): get the exact optimum from full columns
                // first; its band is the narrowest there is
                const int exact = myers_forward<W>(A0, A1, B0, B1, m, nn);
                bd = nw_band(m, nn, 1, 1, exact);
                penalty[p] = exact;
            }
            if (!myers_band_fits(bd)) {
                deferred[atomicAdd(deferred_count, 1u)] = (uint32_t)p;   // nw_wide_kernel, bounded by penalty[p]
            } else if (MODE == kNwCounters) {
                CoverTrace<W> t;
                const int pen = myers_band_cover<W>(A0, A1, B0, B1, m, nn, bd.hi, s_tb + tid, T, t);
                penalty[p] = pen;
                uint32_t P1[2 * W], P2[2 * W], a0[2 * W], a1[2 * W];
                runs3_positions<2 * W>(t.E, t.B, P2);                    // LCM2: read positions under '=' runs >= 3
                const uint32_t g0 = ca.g_off[p];
                const int gn = (int)(ca.g_off[p + 1] - g0);
                if (gops_staged) lcm_toggles<2 * W>(s_gops + (g0 - g_lo), gn, 1, P1);
                else lcm_toggles<2 * W>(ca.g_ops + g0, gn, 1, P1);
                prefix_xor_fill<2 * W>(P1);                              // LCM1: read positions under greedy M runs
#pragma unroll
                for (int w = 0; w < W; ++w) {
                    a0[2 * w] = (uint32_t)A0[w]; a0[2 * w + 1] = (uint32_t)(A0[w] >> 32);
                    a1[2 * w] = (uint32_t)A1[w]; a1[2 * w + 1] = (uint32_t)(A1[w] >> 32);
                }
                const int correct = (ca.answers && ca.answers[p] != INT32_MIN) ? ca.answers[p] : pen;   // benchmark_utils.h:245-248
                c_tot = 1;
                c_nw = pen == correct;
                c_gr = bnd == correct;
                // covers(): when every LCM2 position is an LCM1 position too, each one matches itself at the latest --
                // covered.  Otherwise the sequential subsequence test decides; those pairs (a minority) are queued
                // for cover_hard_kernel, which runs it with full warps instead of a few straggler threads here.
                uint32_t outside = 0u;
#pragma unroll
                for (int w = 0; w < 2 * W; ++w) outside |= P2[w] & ~P1[w];
                if (outside == 0u) {
                    c_cov = 1;
                } else {
                    const uint32_t slot = atomicAdd(ca.hard_count, 1u);
                    uint32_t* h = ca.hard + slot;
#pragma unroll
                    for (int w = 0; w < 2 * W; ++w) {
                        h[(size_t)(4 * w + 0) * ca.hard_cap] = a0[w];
                        h[(size_t)(4 * w + 1) * ca.hard_cap] = a1[w];
                        h[(size_t)(4 * w + 2) * ca.hard_cap] = P1[w];
                        h[(size_t)(4 * w + 3) * ca.hard_cap] = P2[w];
                    }
                    h[(size_t)(8 * W) * ca.hard_cap] = (uint32_t)m;
                }
            } else {
                OpsBackSink sink{MODE == kNwOps ? ops_stage + (size_t)p * ops_stride + ops_stride : nullptr};
                const uint16_t* row_end = sink.wp;
                penalty[p] = myers_band_align<W>(A0, A1, B0, B1, m, nn, bd.hi, TRACE ? s_tb + tid : nullptr, T, sink);
                cnt = (int)(row_end - sink.wp);
            }
        }
        if (MODE == kNwOps) nops_out[p] = (uint32_t)cnt;
    }
    const uint32_t wst = __reduce_or_sync(0xffffffffu, st);
    if (wst && (tid & 31) == 0) atomicOr(status, wst);
    if (MODE == kNwCounters) {
        c_tot = __reduce_add_sync(0xffffffffu, c_tot);
        c_nw = __reduce_add_sync(0xffffffffu, c_nw);
        c_gr = __reduce_add_sync(0xffffffffu, c_gr);
        c_cov = __reduce_add_sync(0xffffffffu, c_cov);
        if ((tid & 31) == 0 && c_tot) {
            atomicAdd(&ca.counters[0], (unsigned long long)c_tot);
            atomicAdd(&ca.counters[1], (unsigned long long)c_nw);
            atomicAdd(&ca.counters[2], (unsigned long long)c_gr);
            atomicAdd(&ca.counters[3], (unsigned long long)c_cov);
        }
    }
}

// The sequential coverage test for the pairs nw_band_kernel (MODE 2) queued: one pair per thread, every lane busy.
template <int W>
__global__ void __launch_bounds__(128)
cover_hard_kernel(const uint32_t* __restrict__ hard, const uint32_t* __restrict__ hard_count, uint32_t hard_cap,
                  unsigned long long* __restrict__ counters) {
    const uint32_t total = *hard_count;
    const uint32_t slot = blockIdx.x * blockDim.x + threadIdx.x;
    if (blockIdx.x * blockDim.x >= total) return;
    uint32_t cov = 0;
    if (slot < total) {
        uint32_t a0[2 * W], a1[2 * W], P1[2 * W], P2[2 * W];
        const uint32_t* h = hard + slot;
#pragma unroll
        for (int w = 0; w < 2 * W; ++w) {
            a0[w] = h[(size_t)(4 * w + 0) * hard_cap];
            a1[w] = h[(size_t)(4 * w + 1) * hard_cap];
            P1[w] = h[(size_t)(4 * w + 2) * hard_cap];
            P2[w] = h[(size_t)(4 * w + 3) * hard_cap];
        }
        cov = (uint32_t)covers_planes<2 * W>(a0, a1, (int)h[(size_t)(8 * W) * hard_cap], P1, P2);
    }
    cov = __reduce_add_sync(0xffffffffu, cov);
    if ((threadIdx.x & 31) == 0 && cov) atomicAdd(&counters[3], (unsigned long long)cov);
}

// Counters of the pairs the banded kernel deferred in MODE 2 (their NW CIGAR is in the staging rows, written back
// to front by nw_wide_kernel).  A handful of pairs per batch at most: one thread each, everything from global memory.
__global__ void __launch_bounds__(128)
nw_deferred_counters_kernel(const uint8_t* __restrict__ reads, const uint32_t* __restrict__ read_off,
                            const int32_t* __restrict__ g_cost, const int32_t* __restrict__ penalty,
                            const uint16_t* __restrict__ n_stage, int ops_stride, const uint32_t* __restrict__ n_cnt,
                            NwCounterArgs ca, const uint32_t* __restrict__ deferred, const uint32_t* __restrict__ deferred_count) {
    const uint32_t total = *deferred_count;
    for (uint32_t it = blockIdx.x * blockDim.x + threadIdx.x; it < total; it += gridDim.x * blockDim.x) {
        const uint32_t p = deferred[it];
        const int pen = penalty[p];
        const int correct = (ca.answers && ca.answers[p] != INT32_MIN) ? ca.answers[p] : pen;
        uint32_t scratch[2 * kCovWords];
        const uint32_t r0 = read_off[p], g0 = ca.g_off[p], cnt = n_cnt[p];
        const int cov = coverage_pair(reads + r0, (int)(read_off[p + 1] - r0), ca.g_ops + g0, (int)(ca.g_off[p + 1] - g0),
                                      n_stage + (size_t)p * ops_stride + (ops_stride - cnt), (int)cnt, scratch, 1);
        atomicAdd(&ca.counters[0], 1ull);
        atomicAdd(&ca.counters[1], (unsigned long long)(pen == correct));
        atomicAdd(&ca.counters[2], (unsigned long long)(g_cost[p] == correct));
        atomicAdd(&ca.counters[3], (unsigned long long)cov);
    }
}

// per-tile sums of an op-count array (input of scan_tiles_kernel)
template <int T>
__global__ void __launch_bounds__(T)
tile_sums_kernel(const uint32_t* __restrict__ nops, long long n, uint32_t* __restrict__ tile_sums) {
    __shared__ uint32_t s_warp[T / 32];
    const int tid = threadIdx.x;
    const long long p = (long long)blockIdx.x * T + tid;
    uint32_t v = p < n ? nops[p] : 0u;
#pragma unroll
    for (int d = 16; d; d >>= 1) v += __shfl_xor_sync(0xffffffffu, v, d);
    if ((tid & 31) == 0) s_warp[tid >> 5] = v;
    __syncthreads();
    if (tid == 0) {
        uint32_t t = 0;
#pragma unroll
        for (int i = 0; i < T / 32; ++i) t += s_warp[i];
        tile_sums[blockIdx.x] = t;
    }
}

}  // namespace gasma
