// metrics.cuh -- K5: the accuracy / coverage counters of the benchmark harness.
//
// Replaces the counter block of benchmark::_run_benchmark (GASMA/benchmark/benchmark_utils.h:238,249-258):
//   total_tests++ ; nw_correct += (nw == correct) ; greedy_correct += (greedy == correct) ;
//   greedy_coverage += covers(LCM(greedy CIGAR, threshold 1), LCM(NW CIGAR, threshold 3))
// with long_consecutive_matching_substring (benchmark_coverage.h:26-67) and covers (:73-91).
// Neither LCM string is materialised: both are streamed from the op lists, and covers() is the
// subsequence test it amounts to (its length pre-check is implied by the scan).
#pragma once
#include "common.cuh"

namespace gasma {

// Streams the characters of long_consecutive_matching_substring(s1, ., CIGAR, threshold).
struct LcmStream {
    const uint16_t* ops;
    int nops, thr;
    const uint8_t* s1;
    int s1_len;
    int oi, rem, s1_index;
    GASMA_HD void init(const uint16_t* ops_, int nops_, int thr_, const uint8_t* s1_, int s1_len_) {
        ops = ops_; nops = nops_; thr = thr_; s1 = s1_; s1_len = s1_len_; oi = 0; rem = 0; s1_index = 0;
    }
    // next LCM character; false when exhausted
    GASMA_HD bool next(uint8_t& c) {
        while (rem == 0) {
            if (oi >= nops) return false;
            const int len = ops[oi] >> 3, code = ops[oi] & 7;
            ++oi;
            if (code == 2) continue;                               // 'D' advances the reference only
            if (code == 1 || code == 4 || len < thr) { s1_index += len; continue; }  // 'I', 'X', short runs
            rem = len;                                             // 'M' / '=' run of length >= threshold
        }
        c = s1_index < s1_len ? s1[s1_index] : (uint8_t)0;         // the reference indexes s1 unchecked (NUL at the end)
        ++s1_index;
        --rem;
        return true;
    }
};

// _check_coverage(s1, s2, greedyCIGAR, nwCIGAR, 1, 3), benchmark_utils.h:214-225
GASMA_HD int coverage_pair(const uint8_t* s1, int s1_len, const uint16_t* g_ops, int g_n, const uint16_t* n_ops, int n_n) {
    LcmStream a, b;
    a.init(g_ops, g_n, 1, s1, s1_len);
    b.init(n_ops, n_n, 3, s1, s1_len);
    uint8_t cb, ca;
    while (b.next(cb)) {
        for (;;) {
            if (!a.next(ca)) return 0;
            if (ca == cb) break;
        }
    }
    return 1;
}

#if defined(__CUDACC__)
// One pair per thread.  The tile's reads are one contiguous byte range of the CSR buffer: it is copied to shared
// memory with coalesced 16-byte loads first, because the coverage test walks each read character by character
// (twice) and would otherwise chase one global byte per step.  `reads_total` = logical size of the reads buffer;
// smem_cap = bytes of dynamic shared memory (a tile that does not fit falls back to global reads).
template <int T>
__global__ void __launch_bounds__(T)
metrics_kernel(const uint8_t* __restrict__ reads, const uint32_t* __restrict__ read_off, long long n,
               const int32_t* __restrict__ g_cost, const uint16_t* __restrict__ g_ops, const uint32_t* __restrict__ g_off,
               const int32_t* __restrict__ nw_pen, const uint16_t* __restrict__ n_ops, const uint32_t* __restrict__ n_off,
               const int32_t* __restrict__ answers, unsigned long long* __restrict__ counters, int smem_cap) {
    extern __shared__ __align__(16) unsigned char s_reads[];
    __shared__ uint32_t s_acc[3];
    const int tid = threadIdx.x;
    if (tid < 3) s_acc[tid] = 0;
    const long long p0 = (long long)blockIdx.x * T;
    const int np = (int)((n - p0) < (long long)T ? (n - p0) : (long long)T);
    const bool cov_wanted = g_ops && n_ops;
    uint32_t base = 0;
    bool staged = false;
    if (cov_wanted) {
        const uint32_t lo = read_off[p0], hi = read_off[p0 + np], total = read_off[n];
        base = lo & ~15u;
        staged = hi - base + 16u <= (uint32_t)smem_cap;   // whole 16-byte chunks are written
        if (staged)
            for (uint32_t a = base + 16u * tid; a < hi; a += 16u * T)
                *reinterpret_cast<uint4*>(s_reads + (a - base)) = load_chunk(reads, a, total);
    }
    __syncthreads();
    const long long p = p0 + tid;
    uint32_t nwc = 0, gc = 0, cov = 0;
    if (p < n) {
        const int nwp = nw_pen[p];
        const int correct = (answers && answers[p] != INT32_MIN) ? answers[p] : nwp;   // benchmark_utils.h:245-248
        nwc = nwp == correct;
        gc = g_cost[p] == correct;
        if (cov_wanted) {
            const uint32_t r0 = read_off[p];
            const uint8_t* s1 = staged ? s_reads + (r0 - base) : reads + r0;
            cov = (uint32_t)coverage_pair(s1, (int)(read_off[p + 1] - r0), g_ops + g_off[p],
                                          (int)(g_off[p + 1] - g_off[p]), n_ops + n_off[p], (int)(n_off[p + 1] - n_off[p]));
        }
    }
    nwc = __reduce_add_sync(0xffffffffu, nwc);
    gc = __reduce_add_sync(0xffffffffu, gc);
    cov = __reduce_add_sync(0xffffffffu, cov);
    if ((tid & 31) == 0) { atomicAdd(&s_acc[0], nwc); atomicAdd(&s_acc[1], gc); atomicAdd(&s_acc[2], cov); }
    __syncthreads();
    if (tid == 0) {
        const long long rem = n - (long long)blockIdx.x * T;
        atomicAdd(&counters[0], (unsigned long long)(rem < T ? rem : T));
        atomicAdd(&counters[1], (unsigned long long)s_acc[0]);
        atomicAdd(&counters[2], (unsigned long long)s_acc[1]);
        atomicAdd(&counters[3], (unsigned long long)s_acc[2]);
    }
}
#endif

}  // namespace gasma
