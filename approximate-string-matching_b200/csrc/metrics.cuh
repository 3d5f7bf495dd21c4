// metrics.cuh -- K5: the accuracy / coverage counters of the benchmark harness.
//
// Replaces the counter block of benchmark::_run_benchmark (GASMA/benchmark/benchmark_utils.h:238,249-258):
//   total_tests++ ; nw_correct += (nw == correct) ; greedy_correct += (greedy == correct) ;
//   greedy_coverage += covers(LCM(greedy CIGAR, threshold 1), LCM(NW CIGAR, threshold 3))
// with long_consecutive_matching_substring (benchmark_coverage.h:26-67) and covers (:73-91).
// Neither LCM string is materialised.  Both are subsets of the READ's positions (an LCM character is s1[index] for
// the indices under qualifying M / '=' runs), so each CIGAR is first turned into a bitmask over read positions
// (two toggles per run + one prefix-XOR fill: no per-character work, no data-dependent inner loops), and covers()
// -- a greedy subsequence test -- becomes ONE loop over the read positions that every thread of a warp runs in
// lock-step.  (The first version streamed characters out of the op lists: correct, but 7 of 32 threads active.)
#pragma once
#include "common.cuh"

namespace gasma {

constexpr int kCovWords = 8;   // read positions covered by the masks: 256 = the NW verifier's max_length

// Positions of long_consecutive_matching_substring(s1, ., CIGAR, threshold) (benchmark_coverage.h:26-67):
// 'D' advances the reference only; 'I', 'X' and runs shorter than the threshold advance the read without
// contributing; 'M' / '=' runs of length >= threshold contribute their read positions.
// mask word w lives at mask[w * stride].  Positions >= 256 are dropped (they lie beyond any read the verifier takes;
// the reference would read the string's NUL there, which never matches an NW character).
GASMA_HD void lcm_positions(const uint16_t* ops, int nops, int thr, uint32_t* mask, int stride) {
    for (int w = 0; w < kCovWords; ++w) mask[w * stride] = 0u;
    int idx = 0;
    for (int oi = 0; oi < nops; ++oi) {
        const int len = ops[oi] >> 3, code = ops[oi] & 7;
        if (code == 2) continue;
        if (!(code == 1 || code == 4 || len < thr)) {   // toggle the run's first position and the one after its last
            if (idx < 32 * kCovWords) mask[(idx >> 5) * stride] ^= 1u << (idx & 31);
            if (idx + len < 32 * kCovWords) mask[((idx + len) >> 5) * stride] ^= 1u << ((idx + len) & 31);
        }
        idx += len;
    }
    uint32_t carry = 0u;   // prefix XOR over all 256 bits: ones between a run's two toggles
    for (int w = 0; w < kCovWords; ++w) {
        uint32_t v = mask[w * stride];
        v ^= v << 1; v ^= v << 2; v ^= v << 4; v ^= v << 8; v ^= v << 16;
        v ^= carry;
        carry = (v >> 31) ? 0xFFFFFFFFu : 0u;
        mask[w * stride] = v;
    }
}

// _check_coverage(s1, s2, greedyCIGAR, nwCIGAR, 1, 3), benchmark_utils.h:214-225; covers() benchmark_coverage.h:73-91.
// scratch: 2 * kCovWords words for this alignment, word i at scratch[i * stride].
GASMA_HD int coverage_pair(const uint8_t* s1, int s1_len, const uint16_t* g_ops, int g_n, const uint16_t* n_ops, int n_n,
                           uint32_t* scratch, int stride) {
    uint32_t* P1 = scratch;
    uint32_t* P2 = scratch + (size_t)kCovWords * stride;
    lcm_positions(g_ops, g_n, 1, P1, stride);
    lcm_positions(n_ops, n_n, 3, P2, stride);
    // j = next NW-LCM position still to be found among the greedy-LCM characters
    int jw = 0;
    uint32_t jword = P2[0];
    while (jword == 0u) {
        if (++jw >= kCovWords) return 1;   // nothing to cover
        jword = P2[jw * stride];
    }
    int j = 32 * jw + ctz64((uint64_t)jword);
    if (j >= s1_len) return 0;             // cannot happen for NW ops of this read (they consume exactly s1_len)
    uint8_t cj = s1[j];
    uint32_t pw = 0u;
    for (int q = 0; q < s1_len && q < 32 * kCovWords; ++q) {
        if ((q & 31) == 0) pw = P1[(q >> 5) * stride];
        if (((pw >> (q & 31)) & 1u) && s1[q] == cj) {
            jword &= jword - 1u;
            while (jword == 0u) {
                if (++jw >= kCovWords) return 1;
                jword = P2[jw * stride];
            }
            j = 32 * jw + ctz64((uint64_t)jword);
            if (j >= s1_len) return 0;
            cj = s1[j];
        }
    }
    return 0;
}

// ---- the same test on 2-bit planes and 32-bit position-mask words (fused into the NW verifier's kernel) -------
// Used by the benchmark body, where the NW walk hands over its '=' steps as bit masks (CoverTrace, nw_myers.cuh) and
// no NW CIGAR is ever written.  Same semantics as coverage_pair: identical characters = identical plane bits
// (inputs outside {A,C,G,T} count as 'A' here, as they do everywhere on the 2-bit path).  NW = number of 32-bit words.
template <int NW>
GASMA_HD uint32_t word32_sel(const uint32_t v[NW], int w) {
    uint32_t x = 0u;
#pragma unroll
    for (int j = 0; j < NW; ++j)
        if (j == w) x = v[j];
    return x;
}
// Read positions under '=' runs of length >= 3, from E (row consumed by '=') and B (a D between this row and the one
// above): p and p+1 are in the same run iff E[p] & E[p+1] & ~B[p]; three in a row make a qualifying triple.
template <int NW>
GASMA_HD void runs3_positions(const uint32_t E[NW], const uint32_t B[NW], uint32_t P2[NW]) {
    uint32_t C[NW];
#pragma unroll
    for (int w = 0; w < NW; ++w) {
        const uint32_t up = (E[w] >> 1) | (w + 1 < NW ? E[w + 1] << 31 : 0u);   // E[p+1] at bit p
        C[w] = E[w] & up & ~B[w];
    }
#pragma unroll
    for (int w = 0; w < NW; ++w) {
        const uint32_t up = (C[w] >> 1) | (w + 1 < NW ? C[w + 1] << 31 : 0u);
        P2[w] = C[w] & up;                                                      // p, p+1, p+2 in one run
    }
    uint32_t c1 = 0u, c2 = 0u;
#pragma unroll
    for (int w = 0; w < NW; ++w) {
        const uint32_t t3 = P2[w];
        P2[w] = t3 | (t3 << 1) | c1 | (t3 << 2) | c2;
        c1 = t3 >> 31;
        c2 = t3 >> 30;
    }
}
// mask of the bits below position n of a 32-bit word; n <= 0 gives 0, n >= 32 all ones
GASMA_HD uint32_t bits_below(int n) {
#if defined(__CUDA_ARCH__)
    return __funnelshift_lc(0xFFFFFFFFu, 0u, (uint32_t)imax(n, 0));   // clamps the shift at 32
#else
    return n <= 0 ? 0u : (n >= 32 ? 0xFFFFFFFFu : (0xFFFFFFFFu >> (32 - n)));
#endif
}
// Read positions under a CIGAR's qualifying runs (see lcm_positions), straight as NW mask words: a run over the
// positions [a, z) sets, in word w, the bits below z - 32w that are not below a - 32w.  Every word is addressed by a
// compile-time index (toggling word a >> 5 of a register array makes the compiler index it at run time, i.e. keep it
// in local memory).  Positions >= 32*NW are dropped.
template <int NW>
GASMA_HD void lcm_mask(const uint16_t* ops, int nops, int thr, uint32_t t[NW]) {
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
    for (int w = 0; w < NW; ++w) t[w] = 0u;
    int idx = 0;
    for (int oi = 0; oi < nops; ++oi) {
        const int len = ops[oi] >> 3, code = ops[oi] & 7;
        if (code == 2) continue;
        if (!(code == 1 || code == 4 || len < thr)) {
            const int a = idx, z = idx + len;
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
            for (int w = 0; w < NW; ++w) t[w] |= bits_below(z - 32 * w) & ~bits_below(a - 32 * w);
        }
        idx += len;
    }
}
GASMA_HD int ctz32m(uint32_t x) {  // x != 0
#if defined(__CUDA_ARCH__)
    return __ffs((int)x) - 1;
#else
    return __builtin_ctz(x);
#endif
}
// covers(): P2's characters (in position order) are a subsequence of P1's characters.  Instead of stepping over
// every read position, each round jumps to the next P1 position that carries the wanted character.
template <int NW>
GASMA_HD int covers_planes(const uint32_t a0[NW], const uint32_t a1[NW], int m, const uint32_t P1[NW], const uint32_t P2[NW]) {
    int jw = 0;
    uint32_t jword = P2[0];
    while (jword == 0u) {
        if (++jw >= NW) return 1;   // nothing to cover
        jword = word32_sel<NW>(P2, jw);
    }
    uint32_t a0j = word32_sel<NW>(a0, jw), a1j = word32_sel<NW>(a1, jw);
    int jb = ctz32m(jword);
    if (32 * jw + jb >= m) return 0;
    uint32_t c0 = 0u - ((a0j >> jb) & 1u), c1 = 0u - ((a1j >> jb) & 1u);   // wanted character, broadcast
#pragma unroll
    for (int w = 0; w < NW; ++w) {
        const int left = m - 32 * w;
        const uint32_t valid = left >= 32 ? 0xFFFFFFFFu : (left <= 0 ? 0u : ((1u << left) - 1u));
        uint32_t prem = P1[w] & valid;
        const uint32_t x0 = a0[w], x1 = a1[w];
        for (;;) {
            const uint32_t cand = prem & ~((x0 ^ c0) | (x1 ^ c1));
            if (cand == 0u) break;
            const uint32_t low = cand & (0u - cand);   // the match: lowest P1 position left with the wanted character
            if (jw == w && low == (1u << jb)) {
                // the wanted position matches itself.  Then so does every directly following position that is in
                // both sets (its own position is the first one left and carries its own character): take the
                // whole run in one round -- '=' runs of the verifier mostly lie inside M runs of the greedy path
                const uint32_t both = (jword & prem) >> jb;
                const uint32_t stop = ~both;
                const int run = stop ? ctz32m(stop) : 32;
                const uint32_t taken = (run >= 32 ? 0xFFFFFFFFu : ((1u << run) - 1u)) << jb;
                jword &= ~taken;
                prem &= ~(taken | (low - 1u));
            } else {
                prem &= ~(low | (low - 1u));     // drop everything up to and including the match
                jword &= jword - 1u;
            }
            while (jword == 0u) {
                if (++jw >= NW) return 1;
                jword = word32_sel<NW>(P2, jw);
                a0j = word32_sel<NW>(a0, jw);
                a1j = word32_sel<NW>(a1, jw);
            }
            jb = ctz32m(jword);
            if (32 * jw + jb >= m) return 0;
            c0 = 0u - ((a0j >> jb) & 1u);
            c1 = 0u - ((a1j >> jb) & 1u);
        }
    }
    return 0;
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
    extern __shared__ __align__(16) unsigned char s_reads[];   // [smem_cap] staged reads, then the position masks
    __shared__ uint32_t s_acc[3];
    uint32_t* s_masks = reinterpret_cast<uint32_t*>(s_reads + smem_cap);   // [2 * kCovWords][T]
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
                                          (int)(g_off[p + 1] - g_off[p]), n_ops + n_off[p], (int)(n_off[p + 1] - n_off[p]),
                                          s_masks + tid, T);
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
