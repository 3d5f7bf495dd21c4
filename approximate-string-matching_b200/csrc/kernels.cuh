// kernels.cuh -- __global__ entry points for K1 (pack2bit) and K2+K3 (greedy), plus the small
// scan/compaction kernels that turn per-pair CIGAR staging rows into one CSR array.
// Built only for sm_100a (see build.py); no other architecture, no CPU path.
#pragma once
#ifndef GASMA_TIMELINE
#define GASMA_TIMELINE 0   // 1: diagnostic build with per-warp time stamps (tools/timeline_probe.py)
#endif
#ifndef GASMA_COOP_TAIL
#define GASMA_COOP_TAIL 4   // open pairs per warp at which the end game takes over (0 = off)
#endif
#include "common.cuh"
#include "pack.cuh"
#include "greedy.cuh"
#include "greedy_coop.cuh"

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

__device__ __forceinline__ uint4 load_chunk(const uint8_t* __restrict__ buf, uint32_t a, uint32_t total) {
    if (a + 16u <= total) return ldg_stream16(buf + a);
    // last partial chunk of the whole buffer: never read past the logical end (rare: kept small, not unrolled).
    // Inlined on purpose: as a real call it breaks the one-chunk-ahead software pipeline of stage_tile (measured:
    // long-scoreboard stalls at the first use of a chunk nearly double).
    uint32_t w[4] = {0u, 0u, 0u, 0u};
#pragma unroll 1
    for (uint32_t i = 0; a + i < total; ++i) {
        const uint32_t b = (uint32_t)buf[a + i] << (8u * (i & 3u));
        if (i < 4u) w[0] |= b; else if (i < 8u) w[1] |= b; else if (i < 12u) w[2] |= b; else w[3] |= b;
    }
    return make_uint4(w[0], w[1], w[2], w[3]);
}

// Packs bytes [base16, end) of `buf` (base16 is 16-byte aligned, total = logical size of buf) into
// stream0/stream1 (one uint16 per chunk).  Returns false if the range does not fit `cap_chunks`.
// Software-pipelined by one chunk (the next 16-byte load is in flight while the current one is packed) and
// deliberately NOT unrolled: the pack body is ~100 instructions and the kernels that inline this are
// instruction-cache sensitive (r01e profile: 90 KB of SASS, "no instruction" the top stall).
__device__ __forceinline__ bool stage_tile(const uint8_t* __restrict__ buf, uint32_t base16, uint32_t end, uint32_t total,
                                           uint16_t* stream0, uint16_t* stream1, int cap_chunks, int tid, int nthreads) {
    int nchunks = (int)((end - base16 + 15u) >> 4);
    bool ok = nchunks <= cap_chunks;
    if (!ok) nchunks = cap_chunks;
    uint4 cur = make_uint4(0u, 0u, 0u, 0u);
    if (tid < nchunks) cur = load_chunk(buf, base16 + 16u * (uint32_t)tid, total);
#pragma unroll 1
    for (int c = tid; c < nchunks; c += nthreads) {
        uint4 nxt = make_uint4(0u, 0u, 0u, 0u);
        if (c + nthreads < nchunks) nxt = load_chunk(buf, base16 + 16u * (uint32_t)(c + nthreads), total);
        uint32_t p0, p1;
        pack_chunk(cur.x, cur.y, cur.z, cur.w, p0, p1);
        stream0[c] = (uint16_t)p0;
        stream1[c] = (uint16_t)p1;
        cur = nxt;
    }
    return ok;
}

// The same for a batch that is ALREADY packed ("2-bit streams", gasma_pack2bit_stream): plane0 / plane1 hold one uint16
// per 16 characters of the CSR buffer, i.e. exactly the layout stage_tile produces, so staging is a copy of the chunk
// range.  Bits that belong to neighbouring strings inside a chunk are masked by the window extraction, as always.
__device__ __forceinline__ bool stage_tile_packed(const uint16_t* __restrict__ plane0, const uint16_t* __restrict__ plane1,
                                                  uint32_t base16, uint32_t end, uint16_t* stream0, uint16_t* stream1,
                                                  int cap_chunks, int tid, int nthreads) {
    int nchunks = (int)((end - base16 + 15u) >> 4);
    const bool ok = nchunks <= cap_chunks;
    if (!ok) nchunks = cap_chunks;
    const uint32_t c0 = base16 >> 4;
#pragma unroll 1
    for (int c = tid; c < nchunks; c += nthreads) {
        stream0[c] = plane0[c0 + c];
        stream1[c] = plane1[c0 + c];
    }
    return ok;
}

// ASCII CSR buffer -> 2-bit streams (K1 for the packed entry points): one 16-byte chunk per thread, HBM-bound
// (reads the bytes once, writes a quarter of them).  `total` = logical size of the buffer.
__global__ void __launch_bounds__(256)
pack2bit_stream_kernel(const uint8_t* __restrict__ seqs, uint32_t total, uint16_t* __restrict__ plane0, uint16_t* __restrict__ plane1) {
    const uint32_t nchunks = (total + 15u) >> 4;
    for (uint32_t c = blockIdx.x * blockDim.x + threadIdx.x; c < nchunks; c += gridDim.x * blockDim.x) {
        const uint4 q = load_chunk(seqs, 16u * c, total);
        uint32_t p0, p1;
        pack_chunk(q.x, q.y, q.z, q.w, p0, p1);
        plane0[c] = (uint16_t)p0;
        plane1[c] = (uint16_t)p1;
    }
}

template <int W>
__host__ __device__ constexpr int stream_chunks(int tile) { return tile * 4 * W + 2; }
// uint16 elements per stream array incl. padding so that extract_plane / pair_init32 may read 2W+3 words from any start
template <int W>
__host__ __device__ constexpr int stream_elems(int tile) { return ((stream_chunks<W>(tile) + 2 * (2 * W + 4) + 7) / 8) * 8; }

// ---------------------------------------------------------------------------------------------
// K2+K3: persistent warps, one alignment per thread, finished threads refilled from a per-warp pool.
//
// The number of _step() calls differs from pair to pair (1 .. >10), so a warp that runs 32 pairs to
// completion idles most of its lanes while the slowest pair finishes (r01a profile: 16.7 of 32 threads active).
// Here every warp owns a POOL of 32 consecutive pairs staged as 2-bit streams in its slice of shared memory;
// a thread whose pair is done takes the next pair of the pool, and when the pool is empty the whole warp
// (all 32 lanes, fully converged) claims the next 32-pair chunk from a global counter and stages it with
// coalesced 16-byte loads.  Results are written by pair index, so they do not depend on the schedule.
// ---------------------------------------------------------------------------------------------
constexpr int kPool = 32;   // pairs per chunk = per warp pool
constexpr int kCsrTileShift = 9, kCsrTile = 1 << kCsrTileShift;   // pairs per CTA of compact_rows_kernel

template <int W>
__host__ __device__ constexpr size_t greedy_smem_bytes(int warps, int nl) {   // nl = lane slots per alignment (KParams::nl)
    return (size_t)nl * (2 * W * 4 + sizeof(typename LaneState<W>::type)) * 32 * warps   // lanes (2W 32-bit words) + state, per thread
           + (size_t)warps * (4 * (size_t)stream_elems<W>(kPool) * 2)   // per-warp 2-bit streams (the pool's offsets live in registers)
           + 64;
}

#if GASMA_TIMELINE
// diagnostic build only (tools/timeline_probe.py): per warp -- start, queue seen dry, last hand-out, exit (globaltimer ns)
__device__ unsigned long long g_tl[4][8192];
__device__ __forceinline__ unsigned long long gtime() { unsigned long long t; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t)); return t; }
#endif
// FLIP: the aligner's mask variant (greedy.cuh:flip_word), 0 = the reference as shipped
template <int W, int WARPS, bool CORR, int FLIP = 0>
__global__ void __launch_bounds__(32 * WARPS, CORR ? 1 : (WARPS == 8 ? 3 : (WARPS == 4 ? (W <= 2 ? 6 : 5) : (WARPS == 2 ? 11 : 1))))
greedy_kernel(KParams P, const uint8_t* __restrict__ reads, const uint32_t* __restrict__ read_off,
              const uint8_t* __restrict__ refs, const uint32_t* __restrict__ ref_off, long long n,
              int32_t* __restrict__ cost, uint16_t* __restrict__ ops_rows, uint16_t* __restrict__ ops_ovf, int ops_stride,
              uint32_t* __restrict__ nops_out, uint32_t* __restrict__ chunk_counter, int refill_threshold,
              uint32_t* __restrict__ status, unsigned long long* __restrict__ step_counter,
              uint32_t* __restrict__ tile_sums,   // optional: op count per kCsrTile pairs, accumulated here (input of compact_rows_kernel)
              // packed input (gasma_greedy_batch_packed): reads / refs then point at the plane-0 streams and these at the
              // plane-1 streams of the two sides; null for ASCII input
              const uint16_t* __restrict__ reads_p1, const uint16_t* __restrict__ refs_p1) {
    constexpr int BITS = 64 * W;
    constexpr int T = 32 * WARPS;
    constexpr int SE = stream_elems<W>(kPool);
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int nl = P.nl;
    using ST = typename LaneState<W>::type;
    ST* s_state = reinterpret_cast<ST*>(smem_raw);
    uint32_t* s_lanes = reinterpret_cast<uint32_t*>(s_state + (size_t)nl * T);
    uint16_t* s_streams = reinterpret_cast<uint16_t*>(s_lanes + (size_t)nl * 2 * W * T);

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    uint16_t* s_r0 = s_streams + (size_t)warp * 4 * SE;
    uint16_t* s_r1 = s_r0 + SE;
    uint16_t* s_f0 = s_r1 + SE;
    uint16_t* s_f1 = s_f0 + SE;
    // zero the stream padding once so every later read is of initialised memory (stale bits past a string's
    // length are masked by extract_plane)
    for (int i = lane; i < SE; i += 32) { s_r0[i] = 0; s_r1[i] = 0; s_f0[i] = 0; s_f1[i] = 0; }
    __syncwarp();

#if GASMA_TIMELINE
    const int wid = (blockIdx.x * WARPS + warp) & 8191;
    if (lane == 0) { g_tl[0][wid] = gtime(); g_tl[1][wid] = 0; g_tl[2][wid] = 0; }
#endif
    const LaneStore<W> S{s_lanes + tid, s_state + tid, T};
    const uint32_t nchunks = (uint32_t)((n + kPool - 1) / kPool);
    const uint32_t rtotal = read_off[n], ftotal = ref_off[n];

    // warp-uniform pool state
    long long pool_p0 = 0;
    int pool_next = 0, pool_count = 0;
    bool pool_ok = true;
    uint32_t rbase = 0, fbase = 0;
    // offsets of the pool's pairs: lane i holds off[p0 + min(i, count)], the end offsets are warp-uniform; a thread
    // that takes pair q reads them with shuffles (the hand-out runs with the warp converged)
    uint32_t pl_ro = 0, pl_fo = 0, pl_re = 0, pl_fe = 0;
    // The NEXT chunk is claimed and brought towards the SM while the current pool is being worked on, one
    // stage per loop iteration so that no stage waits for the previous one's memory round trip:
    //   kClaimed   lane 0 has issued the atomicAdd, its result is not consumed yet
    //   kOffsets   the chunk index is known, the offset loads (lane i: off[p0+i], all lanes: off[p0+cnt]) are issued
    //   kReady     the byte ranges are known, L2 prefetches for both sides are issued
    //   kNone      nothing in flight (start), kDone: the counter ran past the last chunk
    enum { kNone, kClaimed, kOffsets, kReady, kDone };
    int pf = kNone;
    uint32_t nx_chunk = 0;                   // valid in lane 0 from kClaimed, in all lanes from kOffsets
    int nx_count = 0;
    uint32_t nx_ro = 0, nx_fo = 0, nx_re = 0, nx_fe = 0;
    auto advance = [&]() {                   // one pipeline stage; warp-uniform
        if (pf == kNone) {
            if (lane == 0) nx_chunk = atomicAdd(chunk_counter, 1u);
            pf = kClaimed;
        } else if (pf == kClaimed) {
            nx_chunk = __shfl_sync(0xffffffffu, nx_chunk, 0);
            if (nx_chunk >= nchunks) {
                pf = kDone;
#if GASMA_TIMELINE
                if (lane == 0) g_tl[1][wid] = gtime();
#endif
                return;
            }
            const long long p0n = (long long)nx_chunk * kPool;
            nx_count = (int)((n - p0n) < (long long)kPool ? (n - p0n) : (long long)kPool);
            const int i = lane < nx_count ? lane : nx_count;
            nx_ro = read_off[p0n + i];
            nx_fo = ref_off[p0n + i];
            nx_re = read_off[p0n + nx_count];
            nx_fe = ref_off[p0n + nx_count];
            pf = kOffsets;
        } else if (pf == kOffsets) {
            const uint32_t r0 = __shfl_sync(0xffffffffu, nx_ro, 0) & ~127u, f0 = __shfl_sync(0xffffffffu, nx_fo, 0) & ~127u;
#pragma unroll 1
            if (reads_p1 == nullptr) {
#pragma unroll 1
                for (uint32_t a = r0 + 128u * lane; a < nx_re; a += 128u * 32u)
                    asm volatile("prefetch.global.L2 [%0];" ::"l"(reads + a));
#pragma unroll 1
                for (uint32_t a = f0 + 128u * lane; a < nx_fe; a += 128u * 32u)
                    asm volatile("prefetch.global.L2 [%0];" ::"l"(refs + a));
            } else if (lane < 4) {   // packed: 2 bits per character, a chunk's four stream ranges are a line or two each
                const uint8_t* b = lane == 0 ? reads : (lane == 1 ? reinterpret_cast<const uint8_t*>(reads_p1)
                                                                  : (lane == 2 ? refs : reinterpret_cast<const uint8_t*>(refs_p1)));
                const uint32_t lo = (lane < 2 ? r0 : f0) >> 3, hi = (lane < 2 ? nx_re : nx_fe) >> 3;
#pragma unroll 1
                for (uint32_t a = lo & ~127u; a <= hi; a += 128u)
                    asm volatile("prefetch.global.L2 [%0];" ::"l"(b + a));
            }
            pf = kReady;
        }
    };
    // per-thread alignment state
    bool active = false;
    long long my_p = 0;
    PairState ps;
    // a launch that wants no ops (nops_out == nullptr) still stages them in the primary rows (unconditional emit); its
    // "overflow row" is the primary row itself, so longer lists just wrap onto its first slots
    OpSink sink{ops_rows, nops_out ? ops_ovf : ops_rows, nops_out ? ops_stride : kRowOps, 0u, nullptr, 0};
    uint32_t st = 0;
    uint32_t steps_done = 0;

    for (;;) {
        if (pf != kReady && pf != kDone) advance();
        const uint32_t idle = __ballot_sync(0xffffffffu, !active);
        const int n_idle = __popc(idle);
        if (n_idle >= refill_threshold || n_idle == 32) {
            // The claim pipeline advances at the top of every iteration and once more right after a chunk has been
            // staged; with pairs that need one or two steps (low error rates) a pool drains in one or two iterations,
            // so a single call site per iteration falls behind (measured: +25 % kernel time at e = 0.05, +39 % on the
            // Illumina profile).
            if (pool_next == pool_count && pf != kDone) {
                // ---- stage the claimed chunk (whole warp, converged) ----
                // the pool ran empty before the next chunk's pipeline finished (start of the kernel, pairs that finish
                // quickly): finish it now rather than spin through iterations with idle threads
                while (pf != kReady && pf != kDone) advance();
                if (pf == kReady) {
                    pool_p0 = (long long)nx_chunk * kPool;
                    pool_count = nx_count;
                    pool_next = 0;
                    __syncwarp();
                    pl_ro = nx_ro; pl_fo = nx_fo; pl_re = nx_re; pl_fe = nx_fe;
                    rbase = __shfl_sync(0xffffffffu, pl_ro, 0) & ~15u;
                    fbase = __shfl_sync(0xffffffffu, pl_fo, 0) & ~15u;
                    pool_ok = true;
                    if (reads_p1 == nullptr) {
#pragma unroll 1
                        for (int side = 0; side < 2; ++side)   // one instance of the pack code for both sides
                            pool_ok &= stage_tile(side ? refs : reads, side ? fbase : rbase, side ? nx_fe : nx_re,
                                                  side ? ftotal : rtotal, side ? s_f0 : s_r0, side ? s_f1 : s_r1,
                                                  stream_chunks<W>(kPool), lane, 32);
                    } else {
                        pool_ok &= stage_tile_packed(reinterpret_cast<const uint16_t*>(reads), reads_p1, rbase, nx_re, s_r0, s_r1,
                                                     stream_chunks<W>(kPool), lane, 32);
                        pool_ok &= stage_tile_packed(reinterpret_cast<const uint16_t*>(refs), refs_p1, fbase, nx_fe, s_f0, s_f1,
                                                     stream_chunks<W>(kPool), lane, 32);
                    }
                    __syncwarp();
                    pf = kNone;
                    advance();   // claim the chunk after this one right away
                }
            }
            // ---- hand the pool's next pairs to the idle threads ----
            const int q = pool_next + __popc(idle & ((1u << lane) - 1u));
            const int q0 = imin(q, 31), q1 = imin(q + 1, 31);
            const uint32_t ro_q = __shfl_sync(0xffffffffu, pl_ro, q0), fo_q = __shfl_sync(0xffffffffu, pl_fo, q0);
            uint32_t ro_q1 = __shfl_sync(0xffffffffu, pl_ro, q1), fo_q1 = __shfl_sync(0xffffffffu, pl_fo, q1);
            if (q + 1 > 31) { ro_q1 = pl_re; fo_q1 = pl_fe; }
            if (!active && q < pool_count) {
                my_p = pool_p0 + q;
                const int m = (int)(ro_q1 - ro_q);
                const int nn = (int)(fo_q1 - fo_q);
                uint32_t bad = 0;
                if (!pool_ok) bad |= kStatusTile;
                if (m > BITS || nn > BITS) bad |= kStatusLength;
                else if (iabs(nn - m) > kMaxK) bad |= kStatusBand;
                // widened band: the reference's lane arrays end at MAX_K, ours at the nl slots of this context
                else if (CORR && (iabs(nn - m) + P.k > kMaxK || 2 * P.k + 1 + iabs(nn - m) > nl)) bad |= kStatusBand;
                if (bad) {
                    st |= bad;
                    cost[my_p] = INT32_MIN;
                    if (nops_out) nops_out[my_p] = 0u;
                } else {
                    const int rbit = (int)(ro_q - rbase), fbit = (int)(fo_q - fbase);
                    pair_init32<W, CORR>(P, reinterpret_cast<const uint32_t*>(s_r0), reinterpret_cast<const uint32_t*>(s_r1), rbit, m,
                                         reinterpret_cast<const uint32_t*>(s_f0), reinterpret_cast<const uint32_t*>(s_f1), fbit, nn, S, ps);
                    sink.begin((uint32_t)my_p);
                    active = true;
                }
            }
#if GASMA_TIMELINE
            if (lane == 0 && pool_next < pool_count) g_tl[2][wid] = gtime();
#endif
            pool_next = imin(pool_count, pool_next + n_idle);
            __syncwarp();
        }
        if (__ballot_sync(0xffffffffu, active) == 0u) {
            if (pf == kDone && pool_next == pool_count) break;
            continue;   // every pair handed out so far was rejected, or the pool just ran empty
        }
        uint32_t fin_n = 0u, fin_tile = 0u;   // ops of the pair this thread finished in this iteration, and their CSR tile
        if (active) {
            if (pair_step<W, CORR, FLIP>(P, S, ps, sink, st)) {
                pair_finish<W, CORR>(P, S, ps, sink);
                sink.finish();
                cost[my_p] = ps.cost;
                if (nops_out) nops_out[my_p] = (uint32_t)sink.n;
                fin_n = tile_sums ? (uint32_t)sink.n : 0u;
                fin_tile = (uint32_t)(my_p >> kCsrTileShift);
                if (sink.n > ops_stride) st |= kStatusSteps;  // cannot happen: the stride covers the step limit
                steps_done += (uint32_t)ps.steps;
                active = false;
            }
        }
        // Per-tile op counts (input of compact_rows_kernel): the pairs of a pool are consecutive, so the threads that
        // finish in one iteration mostly add to the SAME counter -- 32 same-address atomics per instruction serialise
        // in L2 (measured: +39 % kernel time on the Illumina profile, where a pair takes one or two steps).  The warp
        // groups its finishers by tile and one lane per group adds the group's sum.
        const uint32_t fm = __ballot_sync(0xffffffffu, fin_n != 0u);
        if (fm != 0u && fin_n != 0u) {
            const uint32_t peers = __match_any_sync(fm, fin_tile);
            const uint32_t sum = __reduce_add_sync(peers, fin_n);
            if ((peers & ((1u << lane) - 1u)) == 0u) atomicAdd(tile_sums + fin_tile, sum);
        }
#if GASMA_COOP_TAIL
        // nothing left to hand out and only a few pairs still open in this warp: the end game below finishes them
        if (pf == kDone && pool_next == pool_count && __popc(__ballot_sync(0xffffffffu, active)) <= GASMA_COOP_TAIL) break;
#endif
    }
#if GASMA_COOP_TAIL
    // ---- end game (greedy_coop.cuh): the warp finishes its open pairs one after the other, its threads spread over
    // the pair's lanes; a pair with more than 32 lanes is finished by its owner alone ----
    for (uint32_t todo = __ballot_sync(0xffffffffu, active); todo != 0u; todo &= todo - 1u) {
        const int owner = __ffs((int)todo) - 1;
        const bool fits = CORR ? __shfl_sync(0xffffffffu, ps.ub - ps.lb, owner) < 32 : 2 * P.k + 1 <= 32;
        if (fits) {
            const LaneStore<W> So{s_lanes + (tid - lane + owner), s_state + (tid - lane + owner), T};
            while (!pair_step_coop<W, CORR, FLIP>(P, So, owner, lane, ps, sink, st)) {}
            __syncwarp();
        }
        if (lane == owner) {
            if (!fits)
                while (!pair_step<W, CORR, FLIP>(P, S, ps, sink, st)) {}
            pair_finish<W, CORR>(P, S, ps, sink);
            sink.finish();
            cost[my_p] = ps.cost;
            if (nops_out) nops_out[my_p] = (uint32_t)sink.n;
            if (tile_sums && sink.n) atomicAdd(tile_sums + (uint32_t)(my_p >> kCsrTileShift), (uint32_t)sink.n);
            if (sink.n > ops_stride) st |= kStatusSteps;
            steps_done += (uint32_t)ps.steps;
            active = false;
        }
        __syncwarp();
    }
#endif
#if GASMA_TIMELINE
    if (lane == 0) g_tl[3][wid] = gtime();
#endif
    // one atomic per warp, and only when there is something to report
    const uint32_t wst = __reduce_or_sync(0xffffffffu, st);
    if (wst && lane == 0) atomicOr(status, wst);
    if (step_counter) {
        const uint32_t wsteps = __reduce_add_sync(0xffffffffu, steps_done);
        if (lane == 0) atomicAdd(step_counter, (unsigned long long)wsteps);
    }
}

// ---------------------------------------------------------------------------------------------
// Exclusive scan of per-tile counts (single CTA; a 1 M-pair batch has 7813 tiles).
// base_in (optional, device) is added to every output so several launches can share one CSR array.
// ---------------------------------------------------------------------------------------------
// clear_sums: the sums are zeroed once read (they are accumulated with atomics by the next greedy launch);
// reset_word (optional) is set to 0 at the end (the greedy kernel's chunk counter) -- both save a memset per batch.
__global__ void __launch_bounds__(1024) scan_tiles_kernel(uint32_t* __restrict__ sums, uint32_t* __restrict__ bases,
                                                          int ntiles, uint32_t* total_out, const uint32_t* base_in,
                                                          bool clear_sums = false, uint32_t* reset_word = nullptr) {
    __shared__ uint32_t s_warp[32];
    __shared__ uint32_t s_carry;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid == 0) s_carry = base_in ? *base_in : 0u;   // base_in may be total_out: read here, written at the end
    __syncthreads();
    for (int start = 0; start < ntiles; start += 1024) {
        const int i = start + tid;
        const uint32_t v = i < ntiles ? sums[i] : 0u;
        if (clear_sums && i < ntiles) sums[i] = 0u;
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
    if (tid == 0 && reset_word) *reset_word = 0u;
}

// ---------------------------------------------------------------------------------------------
// CSR of the greedy path in ONE kernel (no separate scan launch): the greedy kernel has already accumulated the op
// count of every kCsrTile-pair tile in `sums`, so a CTA gets its tile's base by adding up its predecessors' counts
// (<= 4 coalesced loads per thread for a 1 M-pair batch, all L2 hits), then scans its own pairs, gathers their ops
// from the 32-byte primary rows (overflow rows for the few longer lists) in shared memory and writes them out 16
// bytes at a time.  op_off[p] = base + tile base + prefix inside the tile; op_off[n] = total.
// Housekeeping for the next batch, so that a greedy step is two launches and no memset:
//   * `sums_next` (the OTHER parity's count array, used by the next greedy launch) is zeroed up to `dirty_next`
//   * `reset_word` (the greedy kernel's chunk counter) is set to 0
//   * base_in / base_out: running op total across the slices of one pipelined host call (two words, alternating,
//     because every CTA reads the base while the last CTA writes the next one)
// ---------------------------------------------------------------------------------------------
constexpr int kCsrSmemOps = kCsrTile * kRowOps;   // a tile whose lists all fit their primary rows always fits

template <int T>
__global__ void __launch_bounds__(T)
compact_rows_kernel(const uint16_t* __restrict__ rows, const uint16_t* __restrict__ ovf, int ops_stride,
                    const uint32_t* __restrict__ nops, const uint32_t* __restrict__ sums, uint32_t* __restrict__ sums_next,
                    int dirty_next, const uint32_t* __restrict__ base_in, uint32_t* __restrict__ base_out, long long n,
                    uint16_t* __restrict__ ops_out, long long ops_cap, uint32_t* __restrict__ op_off,
                    uint32_t* __restrict__ reset_word) {
    static_assert(T == kCsrTile, "one thread per pair of the tile");
    __shared__ uint32_t s_warp[T / 32];
    __shared__ uint32_t s_pref[T / 32];
    __shared__ __align__(16) uint16_t s_ops[kCsrSmemOps + 8];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const long long p = (long long)blockIdx.x * T + tid;
    // housekeeping first: it does not depend on anything below
    for (int i = blockIdx.x * T + tid; i < dirty_next; i += gridDim.x * T) sums_next[i] = 0u;
    if (blockIdx.x == 0 && tid == 0 && reset_word) *reset_word = 0u;
    // ---- base of this tile = ops of all tiles before it ----
    uint32_t acc = 0u;
    for (int i = tid; i < (int)blockIdx.x; i += T) acc += sums[i];
    acc = __reduce_add_sync(0xffffffffu, acc);
    // ---- exclusive scan of this tile's op counts ----
    const uint32_t v = p < n ? nops[p] : 0u;
    uint32_t inc = v;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const uint32_t t = __shfl_up_sync(0xffffffffu, inc, d);
        if (lane >= d) inc += t;
    }
    if (lane == 31) s_warp[warp] = inc;
    if (lane == 0) s_pref[warp] = acc;
    __syncthreads();
    uint32_t wbase = 0u, tile_total = 0u, before = 0u;
#pragma unroll
    for (int i = 0; i < T / 32; ++i) {
        wbase += (i < warp) ? s_warp[i] : 0u;
        tile_total += s_warp[i];
        before += s_pref[i];
    }
    const uint32_t tile_base = (base_in ? *base_in : 0u) + before;
    const uint32_t local = wbase + inc - v;
    const uint32_t off = tile_base + local;
    const uint32_t shift = (uint32_t)((reinterpret_cast<uintptr_t>(ops_out + tile_base) >> 1) & 7u);
    const bool via_smem = tile_total <= (uint32_t)kCsrSmemOps && (long long)tile_base + tile_total <= ops_cap;
    if (p < n) {
        op_off[p] = off;
        if (p == n - 1) {
            op_off[n] = off + v;
            if (base_out) *base_out = off + v;
        }
        if (v <= (uint32_t)kRowKeep) {   // the common case: one 32-byte sector, read as two 16-byte loads
            const uint4* rp = reinterpret_cast<const uint4*>(rows + (size_t)p * kRowPitch);
            uint4 q[2];
            q[0] = v ? rp[0] : make_uint4(0u, 0u, 0u, 0u);
            q[1] = v > 8u ? rp[1] : make_uint4(0u, 0u, 0u, 0u);
            const uint32_t w[8] = {q[0].x, q[0].y, q[0].z, q[0].w, q[1].x, q[1].y, q[1].z, q[1].w};
#pragma unroll
            for (uint32_t e = 0; e < (uint32_t)kRowKeep; ++e) {
                if (e < v) {
                    const uint16_t op = (uint16_t)(w[e >> 1] >> (16u * (e & 1u)));
                    if (via_smem) s_ops[shift + local + e] = op;
                    else if ((long long)(off + e) < ops_cap) ops_out[off + e] = op;
                }
            }
        } else {
            const uint16_t* src = ovf + (size_t)p * ops_stride;
            for (uint32_t e = 0; e < v; ++e) {
                if (via_smem) s_ops[shift + local + e] = src[e];
                else if ((long long)(off + e) < ops_cap) ops_out[off + e] = src[e];
            }
        }
    }
    if (!via_smem) return;   // block-uniform
    __syncthreads();
    // s_ops[shift + k] -> ops_out[tile_base + k]; dst = ops_out + tile_base - shift is 16-byte aligned
    uint16_t* dst = ops_out + tile_base - shift;
    const uint32_t end = shift + tile_total;
    for (uint32_t c = tid; 8u * c < end; c += T) {
        const uint32_t lo = 8u * c;
        if (lo >= shift && lo + 8u <= end) {
            *reinterpret_cast<uint4*>(dst + lo) = *reinterpret_cast<const uint4*>(s_ops + lo);
        } else {
            for (uint32_t e = lo < shift ? shift : lo; e < lo + 8u && e < end; ++e) dst[e] = s_ops[e];
        }
    }
}

// ---------------------------------------------------------------------------------------------
// CSR compaction: op_off[p] = tile_base + exclusive prefix inside the tile; ops copied from the
// per-pair staging rows.  op_off[n] = total.  Writes beyond ops_cap are dropped (status bit set by host).
// ---------------------------------------------------------------------------------------------
constexpr int kCompactSmemOps = 128 * 72;   // ops of one tile gathered in shared memory before the coalesced write-out

// GATHER: collect the tile's ops in shared memory and write them out coalesced (pays off for the NW verifier's
// ~m/3 ops per pair; the greedy path's handful of ops per pair is cheaper written straight from the rows).
template <int T, bool GATHER>
__global__ void __launch_bounds__(T)
compact_ops_kernel(const uint16_t* __restrict__ ops_stage, int ops_stride, const uint32_t* __restrict__ nops,
                   const uint32_t* __restrict__ tile_bases, long long n, uint16_t* __restrict__ ops_out,
                   long long ops_cap, uint32_t* __restrict__ op_off, int from_end) {
    __shared__ uint32_t s_warp[T / 32];
    __shared__ __align__(16) uint16_t s_ops[GATHER ? kCompactSmemOps + 8 : 8];
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
    uint32_t wbase = 0, tile_total = 0;
#pragma unroll
    for (int i = 0; i < T / 32; ++i) {
        wbase += (i < warp) ? s_warp[i] : 0u;
        tile_total += s_warp[i];
    }
    const uint32_t tile_base = tile_bases[blockIdx.x];
    const uint32_t local = wbase + inc - v;      // this pair's first op inside the tile
    const uint32_t off = tile_base + local;
    // The tile's ops are one contiguous range of the output.  They are gathered in shared memory (each thread
    // reads its own row, 8 ops per load) shifted by the output's misalignment, then written out 16 bytes at a time;
    // a tile with more ops than the buffer holds writes straight from the rows.
    const uint32_t shift = (uint32_t)((reinterpret_cast<uintptr_t>(ops_out + tile_base) >> 1) & 7u);
    const bool via_smem = GATHER && tile_total <= (uint32_t)kCompactSmemOps && (long long)tile_base + tile_total <= ops_cap;
    if (p < n) {
        op_off[p] = off;
        if (p == n - 1) op_off[n] = off + v;
        // rows filled back to front (NW traceback) keep their ops at the end of the row.  A row starts on a
        // 16-byte boundary (the stride is a multiple of 8 ops), so it is read 8 ops per load; most pairs need one.
        const uint32_t skip = from_end ? (uint32_t)ops_stride - v : 0u;
        const uint16_t* rowp = ops_stage + (size_t)p * ops_stride;
        for (uint32_t c = skip >> 3; 8u * c < skip + v; ++c) {
            const uint4 q = *reinterpret_cast<const uint4*>(rowp + 8u * c);
            const uint32_t w[4] = {q.x, q.y, q.z, q.w};
#pragma unroll
            for (uint32_t e = 0; e < 8u; ++e) {
                const uint32_t i = 8u * c + e;          // position in the row
                if (i >= skip && i < skip + v) {
                    const uint16_t op = (uint16_t)(w[e >> 1] >> (16u * (e & 1u)));
                    if (via_smem) s_ops[shift + local + i - skip] = op;
                    else if ((long long)(off + i - skip) < ops_cap) ops_out[off + i - skip] = op;
                }
            }
        }
    }
    if (!via_smem) return;   // block-uniform
    __syncthreads();
    // s_ops[shift + k] -> ops_out[tile_base + k]; dst = ops_out + tile_base - shift is 16-byte aligned
    uint16_t* dst = ops_out + tile_base - shift;
    const uint32_t end = shift + tile_total;
    for (uint32_t c = tid; 8u * c < end; c += T) {
        const uint32_t lo = 8u * c;
        if (lo >= shift && lo + 8u <= end) {
            *reinterpret_cast<uint4*>(dst + lo) = *reinterpret_cast<const uint4*>(s_ops + lo);
        } else {
            for (uint32_t e = lo < shift ? shift : lo; e < lo + 8u && e < end; ++e) dst[e] = s_ops[e];
        }
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
