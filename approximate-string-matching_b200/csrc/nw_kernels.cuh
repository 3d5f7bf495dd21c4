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
// the banded kernel adds the per-alignment traceback windows: half of the 64W columns at a time x 2 words x 4 bytes
template <int W>
__host__ __device__ constexpr size_t myers_trace_smem_bytes(int tile) {
    return myers_smem_bytes<W>(tile) + (size_t)32 * W * 2 * 4 * tile;
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

// Banded verifier: one 32-bit window per column (nw_myers.cuh), penalty and -- when TRACE -- the traceback ops in
// one pass.  `bound` is an upper bound on every pair's optimum: the greedy penalty, or the exact optimum from
// nw_myers_kernel (it may then alias `penalty`).  Pairs whose band has more than 32 diagonals are appended to
// `deferred` for nw_wide_kernel (general Gotoh, same traceback rule) with their exact optimum in `penalty` as its bound.
template <int W, int T, bool TRACE>
__global__ void __launch_bounds__(T)
nw_band_kernel(const uint8_t* __restrict__ reads, const uint32_t* __restrict__ read_off, const uint8_t* __restrict__ refs,
               const uint32_t* __restrict__ ref_off, long long n, const int32_t* bound, int32_t* penalty,
               uint16_t* __restrict__ ops_stage, int ops_stride, uint32_t* __restrict__ nops_out,
               uint32_t* __restrict__ deferred, uint32_t* __restrict__ deferred_count, uint32_t* __restrict__ status) {
    constexpr int BITS = 64 * W;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    TraceCol* s_tb = reinterpret_cast<TraceCol*>(smem_raw);                       // [col][T]
    uint16_t* s_r0 = reinterpret_cast<uint16_t*>(s_tb + (TRACE ? (size_t)(BITS / 2) * T : 0));
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
            } else {
                penalty[p] = myers_band_align<W>(A0, A1, B0, B1, m, nn, bd.hi, TRACE ? s_tb + tid : nullptr, T,
                                                 TRACE ? ops_stage + (size_t)p * ops_stride : nullptr, ops_stride, &cnt);
            }
        }
        if (TRACE) nops_out[p] = (uint32_t)cnt;
    }
    const uint32_t wst = __reduce_or_sync(0xffffffffu, st);
    if (wst && (tid & 31) == 0) atomicOr(status, wst);
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
