// kernels.cuh -- __global__ entry points for K1 (pack2bit) and K2+K3 (greedy), plus the small
// scan/compaction kernels that turn per-pair CIGAR staging rows into one CSR array.
// Built only for sm_100a (see build.py); no other architecture, no CPU path.
#pragma once
#include "common.cuh"
#include "pack.cuh"
#include "greedy.cuh"

namespace gasma {

// ---------------------------------------------------------------------------------------------
// Tile staging: a CTA owns `np` consecutive strings of one CSR buffer = one contiguous byte range.
// Threads sweep that range in 16-byte chunks (coalesced, streaming loads), pack every chunk to
// 16+16 plane bits and store them in shared memory as a bit stream.
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ uint4 ldg_stream16(const void* p) {
    uint4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
    return r;
}

// Packs bytes [base16, end) of `buf` (base16 is 16-byte aligned, total = logical size of buf) into
// stream0/stream1 (one uint16 per chunk).  Returns false if the range does not fit `cap_chunks`.
__device__ __forceinline__ bool stage_tile(const uint8_t* __restrict__ buf, uint32_t base16, uint32_t end, uint32_t total,
                                           uint16_t* stream0, uint16_t* stream1, int cap_chunks, int tid, int nthreads) {
    int nchunks = (int)((end - base16 + 15u) >> 4);
    bool ok = nchunks <= cap_chunks;
    if (!ok) nchunks = cap_chunks;
    for (int c = tid; c < nchunks; c += nthreads) {
        const uint32_t a = base16 + 16u * (uint32_t)c;
        uint4 v;
        if (a + 16u <= total) {
            v = ldg_stream16(buf + a);
        } else {  // last partial chunk of the whole buffer: never read past the logical end
            uint32_t w[4] = {0u, 0u, 0u, 0u};
            for (uint32_t i = 0; a + i < total && i < 16u; ++i) w[i >> 2] |= (uint32_t)buf[a + i] << (8u * (i & 3u));
            v = make_uint4(w[0], w[1], w[2], w[3]);
        }
        uint32_t p0, p1;
        pack_chunk(v.x, v.y, v.z, v.w, p0, p1);
        stream0[c] = (uint16_t)p0;
        stream1[c] = (uint16_t)p1;
    }
    return ok;
}

template <int W>
__host__ __device__ constexpr int stream_chunks(int tile) { return tile * 4 * W + 2; }
// uint16 elements per stream array incl. padding so that extract_plane may read 2W+1 words past any start
template <int W>
__host__ __device__ constexpr int stream_elems(int tile) { return ((stream_chunks<W>(tile) + 2 * (2 * W + 2) + 7) / 8) * 8; }

template <int W>
__host__ __device__ constexpr size_t greedy_smem_bytes(int tile, int k) {
    return (size_t)(2 * k + 1) * (2 * W + 1) * 8 * tile   // lanes + state
           + 4 * (size_t)stream_elems<W>(tile) * 2        // read/ref x plane0/plane1 streams
           + 2 * (size_t)(tile + 1) * 4                   // offsets
           + 64;
}

// ---------------------------------------------------------------------------------------------
// K2+K3: one CTA per tile of T pairs, one thread per pair.
// ---------------------------------------------------------------------------------------------
template <int W, int T>
__global__ void __launch_bounds__(T)
greedy_kernel(KParams P, const uint8_t* __restrict__ reads, const uint32_t* __restrict__ read_off,
              const uint8_t* __restrict__ refs, const uint32_t* __restrict__ ref_off, long long n,
              int32_t* __restrict__ cost, uint16_t* __restrict__ ops_stage, int ops_stride,
              uint32_t* __restrict__ nops_out, uint32_t* __restrict__ tile_sums,
              uint32_t* __restrict__ status, unsigned long long* __restrict__ step_counter) {
    constexpr int BITS = 64 * W;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int nl = 2 * P.k + 1;
    uint64_t* s_lanes = reinterpret_cast<uint64_t*>(smem_raw);
    uint64_t* s_state = s_lanes + (size_t)nl * 2 * W * T;
    uint16_t* s_r0 = reinterpret_cast<uint16_t*>(s_state + (size_t)nl * T);
    uint16_t* s_r1 = s_r0 + stream_elems<W>(T);
    uint16_t* s_f0 = s_r1 + stream_elems<W>(T);
    uint16_t* s_f1 = s_f0 + stream_elems<W>(T);
    uint32_t* s_roff = reinterpret_cast<uint32_t*>(s_f1 + stream_elems<W>(T));
    uint32_t* s_foff = s_roff + (T + 1);
    __shared__ uint32_t s_warp_sums[T / 32];

    const int tid = threadIdx.x;
    const long long p0 = (long long)blockIdx.x * T;
    const int np = (int)((n - p0) < (long long)T ? (n - p0) : (long long)T);

    for (int i = tid; i <= np; i += T) {
        s_roff[i] = read_off[p0 + i];
        s_foff[i] = ref_off[p0 + i];
    }
    // zero the stream padding so every later read is of initialised memory
    for (int i = tid; i < stream_elems<W>(T); i += T) { s_r0[i] = 0; s_r1[i] = 0; s_f0[i] = 0; s_f1[i] = 0; }
    __syncthreads();
    const uint32_t rbase = s_roff[0] & ~15u, fbase = s_foff[0] & ~15u;
    const uint32_t rtotal = read_off[n], ftotal = ref_off[n];
    bool ok = stage_tile(reads, rbase, s_roff[np], rtotal, s_r0, s_r1, stream_chunks<W>(T), tid, T);
    ok &= stage_tile(refs, fbase, s_foff[np], ftotal, s_f0, s_f1, stream_chunks<W>(T), tid, T);
    __syncthreads();

    uint32_t my_nops = 0, st = 0;
    int steps = 0;
    if (tid < np) {
        const int m = (int)(s_roff[tid + 1] - s_roff[tid]);
        const int nn = (int)(s_foff[tid + 1] - s_foff[tid]);
        if (!ok) st |= kStatusTile;
        if (m > BITS || nn > BITS) st |= kStatusLength;
        else if (iabs(nn - m) > kMaxK) st |= kStatusBand;
        int c = INT32_MIN;
        if (st == 0) {
            uint64_t A0[W], A1[W], B0[W], B1[W];
            const int rbit = (int)(s_roff[tid] - rbase), fbit = (int)(s_foff[tid] - fbase);
            extract_plane<W>(reinterpret_cast<const uint32_t*>(s_r0), rbit, m, A0);
            extract_plane<W>(reinterpret_cast<const uint32_t*>(s_r1), rbit, m, A1);
            extract_plane<W>(reinterpret_cast<const uint32_t*>(s_f0), fbit, nn, B0);
            extract_plane<W>(reinterpret_cast<const uint32_t*>(s_f1), fbit, nn, B1);
            LaneStore S{s_lanes + tid, s_state + tid, T};
            OpSink sink{ops_stage ? ops_stage + (size_t)(p0 + tid) * ops_stride : nullptr, ops_stride, 0};
            st = greedy_pair<W>(P, A0, A1, B0, B1, m, nn, S, sink, c, steps);
            my_nops = (uint32_t)sink.n;
            if (sink.n > ops_stride) st |= kStatusSteps;  // cannot happen: stride covers the step limit
        }
        cost[p0 + tid] = c;
        if (nops_out) nops_out[p0 + tid] = my_nops;
    }
    // one atomic per warp, and only when there is something to report
    const uint32_t wst = __reduce_or_sync(0xffffffffu, st);
    if (wst && (tid & 31) == 0) atomicOr(status, wst);
    if (step_counter) {
        const uint32_t wsteps = __reduce_add_sync(0xffffffffu, (uint32_t)steps);
        if ((tid & 31) == 0) atomicAdd(step_counter, (unsigned long long)wsteps);
    }
    if (tile_sums) {  // per-tile op count for the CSR scan
        uint32_t v = my_nops;
#pragma unroll
        for (int d = 16; d; d >>= 1) v += __shfl_xor_sync(0xffffffffu, v, d);
        if ((tid & 31) == 0) s_warp_sums[tid >> 5] = v;
        __syncthreads();
        if (tid == 0) {
            uint32_t t = 0;
#pragma unroll
            for (int i = 0; i < T / 32; ++i) t += s_warp_sums[i];
            tile_sums[blockIdx.x] = t;
        }
    }
}

// ---------------------------------------------------------------------------------------------
// Exclusive scan of per-tile counts (single CTA; a 1 M-pair batch has 7813 tiles).
// base_in (optional, device) is added to every output so several launches can share one CSR array.
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(1024) scan_tiles_kernel(const uint32_t* __restrict__ sums, uint32_t* __restrict__ bases,
                                                          int ntiles, uint32_t* __restrict__ total_out) {
    __shared__ uint32_t s_warp[32];
    __shared__ uint32_t s_carry;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid == 0) s_carry = 0;
    __syncthreads();
    for (int start = 0; start < ntiles; start += 1024) {
        const int i = start + tid;
        const uint32_t v = i < ntiles ? sums[i] : 0u;
        uint32_t inc = v;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            uint32_t t = __shfl_up_sync(0xffffffffu, inc, d);
            if (lane >= d) inc += t;
        }
        if (lane == 31) s_warp[warp] = inc;
        __syncthreads();
        if (warp == 0) {
            uint32_t w = s_warp[lane], winc = w;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                uint32_t t = __shfl_up_sync(0xffffffffu, winc, d);
                if (lane >= d) winc += t;
            }
            s_warp[lane] = winc - w;  // exclusive prefix of warp totals
        }
        __syncthreads();
        const uint32_t carry = s_carry;
        const uint32_t excl = carry + s_warp[warp] + inc - v;
        if (i < ntiles) bases[i] = excl;
        __syncthreads();
        if (tid == 1023) s_carry = excl + v;
        __syncthreads();
    }
    if (tid == 0 && total_out) *total_out = s_carry;
}

// ---------------------------------------------------------------------------------------------
// CSR compaction: op_off[p] = tile_base + exclusive prefix inside the tile; ops copied from the
// per-pair staging rows.  op_off[n] = total.  Writes beyond ops_cap are dropped (status bit set by host).
// ---------------------------------------------------------------------------------------------
template <int T>
__global__ void __launch_bounds__(T)
compact_ops_kernel(const uint16_t* __restrict__ ops_stage, int ops_stride, const uint32_t* __restrict__ nops,
                   const uint32_t* __restrict__ tile_bases, long long n, uint16_t* __restrict__ ops_out,
                   long long ops_cap, uint32_t* __restrict__ op_off, int from_end) {
    __shared__ uint32_t s_warp[T / 32];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const long long p = (long long)blockIdx.x * T + tid;
    const uint32_t v = p < n ? nops[p] : 0u;
    uint32_t inc = v;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        uint32_t t = __shfl_up_sync(0xffffffffu, inc, d);
        if (lane >= d) inc += t;
    }
    if (lane == 31) s_warp[warp] = inc;
    __syncthreads();
    uint32_t wbase = 0;
#pragma unroll
    for (int i = 0; i < T / 32; ++i) wbase += (i < warp) ? s_warp[i] : 0u;
    const uint32_t off = tile_bases[blockIdx.x] + wbase + inc - v;
    if (p < n) {
        op_off[p] = off;
        if (p == n - 1) op_off[n] = off + v;
        // rows filled back to front (NW traceback) keep their ops at the end of the row
        const uint16_t* src = ops_stage + (size_t)p * ops_stride + (from_end ? (uint32_t)ops_stride - v : 0u);
        for (uint32_t i = 0; i < v; ++i)
            if ((long long)(off + i) < ops_cap) ops_out[off + i] = src[i];
    }
}

// ---------------------------------------------------------------------------------------------
// K1 standalone: plane0/plane1[i*W + w] for a batch of strings (sse3_convert2bit1 per string).
// ---------------------------------------------------------------------------------------------
template <int W, int T>
__global__ void __launch_bounds__(T)
pack2bit_kernel(const uint8_t* __restrict__ seqs, const uint32_t* __restrict__ off, long long n,
                uint64_t* __restrict__ plane0, uint64_t* __restrict__ plane1, uint32_t* __restrict__ status) {
    constexpr int BITS = 64 * W;
    __shared__ __align__(16) uint16_t s_0[stream_elems<W>(T)];
    __shared__ __align__(16) uint16_t s_1[stream_elems<W>(T)];
    __shared__ uint32_t s_off[T + 1];
    const int tid = threadIdx.x;
    const long long p0 = (long long)blockIdx.x * T;
    const int np = (int)((n - p0) < (long long)T ? (n - p0) : (long long)T);
    for (int i = tid; i <= np; i += T) s_off[i] = off[p0 + i];
    for (int i = tid; i < stream_elems<W>(T); i += T) { s_0[i] = 0; s_1[i] = 0; }
    __syncthreads();
    const uint32_t base = s_off[0] & ~15u;
    bool ok = stage_tile(seqs, base, s_off[np], off[n], s_0, s_1, stream_chunks<W>(T), tid, T);
    __syncthreads();
    if (tid < np) {
        const int len = (int)(s_off[tid + 1] - s_off[tid]);
        uint64_t a[W], b[W];
        if (!ok || len > BITS) {
            atomicOr(status, kStatusLength | (ok ? 0u : kStatusTile));
#pragma unroll
            for (int j = 0; j < W; ++j) a[j] = b[j] = 0ull;
        } else {
            const int bit = (int)(s_off[tid] - base);
            extract_plane<W>(reinterpret_cast<const uint32_t*>(s_0), bit, len, a);
            extract_plane<W>(reinterpret_cast<const uint32_t*>(s_1), bit, len, b);
        }
#pragma unroll
        for (int j = 0; j < W; ++j) {
            plane0[(p0 + tid) * W + j] = a[j];
            plane1[(p0 + tid) * W + j] = b[j];
        }
    }
}

}  // namespace gasma
