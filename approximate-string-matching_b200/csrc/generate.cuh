// generate.cuh -- synthetic (read, ref) pairs on the device.
//
// Restates Dataset (GASMA/benchmark/benchmark_dataset.h:100-186): read = L iid uniform ACGT (:100-110);
// ref = copy of the read, then num_errors edits applied sequentially to the ref (:147-186): with
// probability mismatch_rate overwrite a random position with a random base (may equal the old one,
// :113-120), otherwise delete (:121-132) or insert (:133-146) at a random position, 50/50 (:169).
// libc rand() (seeded with time(), :190,:223) is replaced by a counter RNG -- a splitmix64 stream keyed by
// (seed, pair index) -- so that any GPU regenerates pair i identically and the CPU baseline can regenerate
// the very same pairs.  Mode 2 is a per-base independent profile (README.md:71-76, SRR611076-shaped).
// The CPU twin used by the tests lives in oracle/gasma_oracle.c (go_generate_pair).
#pragma once
#include "common.cuh"

namespace gasma {

struct GenCfg {
    uint64_t seed;
    int length, mode, num_errors;
    uint64_t thr_mismatch, thr_sub, thr_ins, thr_del;
};

struct Rng {
    uint64_t s;
    GASMA_HD uint64_t next() {
        uint64_t z = (s += 0x9E3779B97F4A7C15ull);
        z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
        z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
        return z ^ (z >> 31);
    }
    GASMA_HD uint32_t u32() { return (uint32_t)(next() >> 32); }
    GASMA_HD uint32_t below(uint32_t n) { return (uint32_t)(((uint64_t)u32() * n) >> 32); }
    static GASMA_HD uint64_t mix(uint64_t z) {
        z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
        z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
        return z ^ (z >> 31);
    }
    // The stream of a pair starts at a HASHED point of the 2^64 cycle.  (Round 1 started pair p at seed*A + p*G + C with
    // G the generator's own increment: pair p+1's stream was pair p's stream shifted by one draw, so neighbouring
    // reads were shifted copies of each other -- not the independent simulated reads the workload is meant to be; it
    // showed as +-9 % kernel time between batches of identical total work.)
    GASMA_HD void seed(uint64_t seed_, uint64_t pair) {
        s = mix(mix(seed_ + 0x2545F4914F6CDD1Dull) ^ (pair * 0xD1B54A32D192ED03ull + 0x9E3779B97F4A7C15ull));
    }
};

GASMA_HD int gen_ref_capacity(const GenCfg& c) { return c.mode == 2 ? 2 * c.length + 8 : c.length + c.num_errors + 1; }

GASMA_HD char gen_base(uint32_t r) { return (char)((0x54474341u >> (8u * (r >> 30))) & 0xFFu); }  // "ACGT"

// Generates pair `pair`.  With WRITE=false only the RNG is replayed and the ref length returned.
template <bool WRITE>
GASMA_HD int gen_pair(const GenCfg& c, uint64_t pair, char* read, char* ref) {
    Rng r;
    r.seed(c.seed, pair);
    const int L = c.length;
    if (c.mode == 2) {
        // the read is needed to derive the ref, so draw bases one at a time (same draw order as the oracle:
        // all L read draws first, then the per-base draws)
        Rng rr = r;  // second cursor positioned after the L read draws
        for (int i = 0; i < L; ++i) (void)rr.next();
        int len = 0;
        for (int i = 0; i < L; ++i) {
            const char base = gen_base(r.u32());
            if (WRITE) read[i] = base;
            const uint64_t u = rr.u32();
            if (u < c.thr_del) continue;
            if (u < c.thr_ins) {
                const char ins = gen_base(rr.u32());
                if (WRITE) { ref[len] = ins; ref[len + 1] = base; }
                len += 2;
                continue;
            }
            if (u < c.thr_sub) {
                const int b = base == 'A' ? 0 : base == 'C' ? 1 : base == 'G' ? 2 : 3;
                const uint32_t sub = (uint32_t)(b + 1 + (int)rr.below(3)) & 3u;
                if (WRITE) ref[len] = gen_base(sub << 30);
                ++len;
                continue;
            }
            if (WRITE) ref[len] = base;
            ++len;
        }
        return len;
    }
    for (int i = 0; i < L; ++i) {
        const char base = gen_base(r.u32());
        if (WRITE) { read[i] = base; ref[i] = base; }
    }
    int len = L;
    int ne = c.num_errors;
    if (c.mode == 1) ne = c.num_errors > 0 ? (int)r.below((uint32_t)c.num_errors) : 0;
    for (int t = 0; t < ne; ++t) {
        const uint64_t u = r.u32();
        if (u < c.thr_mismatch) {
            if (len == 0) continue;
            const int pos = (int)r.below((uint32_t)len);
            const char b = gen_base(r.u32());
            if (WRITE) ref[pos] = b;
        } else if (r.below(2) == 0) {  // deletion
            if (len == 0) continue;
            const int pos = (int)r.below((uint32_t)len);
            if (WRITE) for (int i = pos; i < len - 1; ++i) ref[i] = ref[i + 1];
            --len;
        } else {                        // insertion
            const int pos = len ? (int)r.below((uint32_t)len) : 0;
            if (WRITE) for (int i = len; i > pos; --i) ref[i] = ref[i - 1];
            ++len;
            const char b = gen_base(r.u32());
            if (WRITE) ref[pos] = b;
        }
    }
    return len;
}

#if defined(__CUDACC__)
constexpr int kGenMaxRef = 2 * 1024 + 8 + 8;

// pass 1: ref lengths (+ per-tile sums for the CSR scan); read offsets are i*L
template <int T>
__global__ void __launch_bounds__(T)
gen_lengths_kernel(GenCfg c, unsigned long long first, long long n, uint32_t* __restrict__ ref_len,
                   uint32_t* __restrict__ tile_sums, uint32_t* __restrict__ read_off) {
    __shared__ uint32_t s_warp[T / 32];
    const int tid = threadIdx.x;
    const long long p = (long long)blockIdx.x * T + tid;
    uint32_t v = 0;
    if (p < n) {
        v = (uint32_t)gen_pair<false>(c, first + (unsigned long long)p, nullptr, nullptr);
        ref_len[p] = v;
        read_off[p] = (uint32_t)(p * c.length);
        if (p == n - 1) read_off[n] = (uint32_t)(n * c.length);
    }
    uint32_t s = v;
#pragma unroll
    for (int d = 16; d; d >>= 1) s += __shfl_xor_sync(0xffffffffu, s, d);
    if ((tid & 31) == 0) s_warp[tid >> 5] = s;
    __syncthreads();
    if (tid == 0) {
        uint32_t t = 0;
        for (int i = 0; i < T / 32; ++i) t += s_warp[i];
        tile_sums[blockIdx.x] = t;
    }
}

// pass 2: offsets from tile bases + in-tile scan, then the bases themselves
template <int T>
__global__ void __launch_bounds__(T)
gen_write_kernel(GenCfg c, unsigned long long first, long long n, const uint32_t* __restrict__ ref_len,
                 const uint32_t* __restrict__ tile_bases, char* __restrict__ reads, char* __restrict__ refs,
                 uint32_t* __restrict__ ref_off) {
    __shared__ uint32_t s_warp[T / 32];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const long long p = (long long)blockIdx.x * T + tid;
    const uint32_t v = p < n ? ref_len[p] : 0u;
    uint32_t inc = v;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        uint32_t t = __shfl_up_sync(0xffffffffu, inc, d);
        if (lane >= d) inc += t;
    }
    if (lane == 31) s_warp[warp] = inc;
    __syncthreads();
    uint32_t wbase = 0;
    for (int i = 0; i < T / 32; ++i) wbase += (i < warp) ? s_warp[i] : 0u;
    const uint32_t off = tile_bases[blockIdx.x] + wbase + inc - v;
    if (p < n) {
        ref_off[p] = off;
        if (p == n - 1) ref_off[n] = off + v;
        char rbuf[1024 + 8];
        char fbuf[kGenMaxRef];
        const int len = gen_pair<true>(c, first + (unsigned long long)p, rbuf, fbuf);
        char* rd = reads + (size_t)p * c.length;
        for (int i = 0; i < c.length; ++i) rd[i] = rbuf[i];
        char* fd = refs + off;
        for (int i = 0; i < len; ++i) fd[i] = fbuf[i];
    }
}
#endif

}  // namespace gasma
