// common.cuh -- shared definitions for the sm_100a kernels of libgasma_b200.so.
//
// Everything that is pure per-pair arithmetic is written as GASMA_HD (host+device) so the very
// same code can be exercised on the CPU by tests/host_sim.cpp against the oracle while there is no
// GPU in the build container.  That harness is test infrastructure; the shipped library only ever
// runs these functions inside kernels.
#pragma once
#include <stdint.h>

#if defined(__CUDACC__)
#define GASMA_HD __host__ __device__ __forceinline__
#define GASMA_D __device__ __forceinline__
#else
#define GASMA_HD inline
#define GASMA_D inline
#endif

namespace gasma {

constexpr int kMaxK = 50;          // hurdle_matrix.h:8
constexpr int kTile = 128;         // pairs per CTA tile (= threads per CTA in the thread-per-pair kernels)

// device-side status bits (OR-ed into ctx->d_status by kernels, read back by gasma_synchronize)
enum : uint32_t {
    kStatusLength = 1u,   // a string longer than 64*W
    kStatusBand = 2u,     // |n-m| > MAX_K
    kStatusSteps = 4u,    // traversal did not terminate within the step limit
    kStatusTile = 8u,     // tile byte range larger than the shared-memory stream (implies kStatusLength)
};

// Parameters of one aligner configuration, passed by value to kernels.
// match/mismatch/indel significances are computed on the HOST with libm's log() exactly as
// hurdle_matrix.h:536-538 does, so the device constants are bit-identical to the reference's.
struct KParams {
    int k, x, o, e;
    int global;          // 1 = GLOBAL (utils.h:554)
    double ms, mms, ids;
    // hurdle_matrix.h:501-512, :634-641 (-DCORRECTION build of the reference): band widened towards lane n-m.
    // nl = lane slots kept per alignment: 2k+1, or 2k+1+max_delta when correction is on (pairs with a larger
    // |n-m| are flagged kStatusBand)
    int correction = 0;
    int nl = 0;
};

// ---- clamped 64-bit shifts: amount >= 64 gives 0 (PTX shl/shr.b64 semantics; C++ << would be UB) ----
GASMA_HD uint64_t shl64(uint64_t x, int s) {
#if defined(__CUDA_ARCH__)
    uint64_t r;
    asm("shl.b64 %0, %1, %2;" : "=l"(r) : "l"(x), "r"(s));
    return r;
#else
    return (unsigned)s >= 64u ? 0ull : x << s;
#endif
}
GASMA_HD uint64_t shr64(uint64_t x, int s) {
#if defined(__CUDA_ARCH__)
    uint64_t r;
    asm("shr.b64 %0, %1, %2;" : "=l"(r) : "l"(x), "r"(s));
    return r;
#else
    return (unsigned)s >= 64u ? 0ull : x >> s;
#endif
}
GASMA_HD int popc64(uint64_t x) {
#if defined(__CUDA_ARCH__)
    return __popcll(x);
#else
    return __builtin_popcountll(x);
#endif
}
GASMA_HD int ctz64(uint64_t x) {  // x != 0
#if defined(__CUDA_ARCH__)
    return __ffsll((long long)x) - 1;
#else
    return __builtin_ctzll(x);
#endif
}
GASMA_HD int imin(int a, int b) { return a < b ? a : b; }
GASMA_HD int imax(int a, int b) { return a > b ? a : b; }
GASMA_HD int iabs(int a) { return a < 0 ? -a : a; }

// all-ones from bit s upwards, for word j of a multi-word vector (s may be any int)
GASMA_HD uint64_t ones_from(int s, int j) { return shl64(~0ull, imax(s - 64 * j, 0)); }

// ---- lane geometry -------------------------------------------------------------------------
// switch_lane_penalty, utils.h:576-579
// = min(d,1)*(o-e) + e*d with d = |l1-l2|: branch-free, two multiply-adds
GASMA_HD int slp(int l1, int l2, int o, int e) {
    const int d = iabs(l1 - l2);
    return imin(d, 1) * (o - e) + e * d;
}
// switch_forward_column, utils.h:587-593
// Lanes on opposite sides of lane 0 (sign bits differ, neither is 0) cost |l1|, otherwise max(|l1|-|l2|, 0).  When
// one of them is 0 both forms agree, so the sign test alone decides: max(|l1| - (opposite ? 0 : |l2|), 0).
GASMA_HD int sfc(int l1, int l2) {
    const int a2 = ((l1 ^ l2) < 0) ? 0 : iabs(l2);
    return imax(iabs(l1) - a2, 0);
}
// highways::_calculate_destination, hurdle_matrix.h:58-68, closed form: min(m+min(l,0), n-max(l,0))
GASMA_HD int lane_destination(int m, int n, int l) { return imin(m + imin(l, 0), n - imax(l, 0)); }

// FP64 without contraction: the reference is pinned to -ffp-contract=off (SURVEY.md hazard 2)
GASMA_HD double dmul(double a, double b) {
#if defined(__CUDA_ARCH__)
    return __dmul_rn(a, b);
#else
    volatile double r = a * b;
    return r;
#endif
}
GASMA_HD double dadd(double a, double b) {
#if defined(__CUDA_ARCH__)
    return __dadd_rn(a, b);
#else
    volatile double r = a + b;
    return r;
#endif
}

}  // namespace gasma
