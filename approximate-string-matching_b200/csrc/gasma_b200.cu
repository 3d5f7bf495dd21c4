// gasma_b200.cu -- the C ABI of libgasma_b200.so (include/gasma_b200.h) on top of the sm_100a kernels.
// Plain CUDA runtime; no torch types, no CPU compute path: every result is produced by a kernel.
#include "../../include/gasma_b200.h"

#include <cuda_runtime.h>
#include <math.h>
#include <stdarg.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <chrono>
#include <string>

#include "kernels.cuh"
#include "leap.cuh"
#include "greedy_warp.cuh"
#include "nw_kernels.cuh"
#include "generate.cuh"

using namespace gasma;

namespace {

thread_local std::string g_create_error;

struct DevBuf {
    void* p = nullptr;
    size_t cap = 0;
};

}  // namespace

struct gasma_async_state;   // async_entry.inl: the two in-flight slots of gasma_greedy_submit / gasma_greedy_wait

struct gasma_ctx {
    gasma_async_state* async = nullptr;
    gasma_params params;
    KParams kp;
    int W = 2;
    int device = 0;
    int sm_count = 148;
    int refill_threshold = 8;   // idle threads in a warp that trigger a refill from the pool (greedy kernel)
    int greedy_warps = 0;       // warps per CTA of the thread-per-pair greedy kernel: 0 = most resident warps (env GASMA_GREEDY_WARPS: 2, 4, 8)
    bool counter_clean = false;  // the greedy kernel's chunk counter is 0 (left so by scan_tiles_kernel of the previous batch)
    int auto_warps = 0;         // the choice made for the current (W, lanes); reset by derive()
    // packed input of the launch being enqueued (gasma_greedy_batch_packed): plane-1 streams of reads / refs; the
    // plane-0 streams travel in the reads / refs arguments.  Null = ASCII input.  Set and cleared by the packed entry points.
    const uint16_t* pk_r1 = nullptr;
    const uint16_t* pk_f1 = nullptr;
    DevBuf nccl_words;          // scratch of gasma_allreduce_counters_host
    DevBuf d_r1, d_f1;          // device mirrors of the plane-1 streams (host-pointer packed call)
    int mask_variant = 0;       // gasma_set_mask_variant: 0 flip_short_hurdles(1), 1 flip_short_hurdles(2), 2 + flip_short_matches(1)
    int correction = 0;         // gasma_set_correction: widened band of the reference's -DCORRECTION build
    int max_delta = 0;          // largest |ref_len - read_len| the widened band has lane slots for
    cudaStream_t stream = nullptr;      // stream all work goes to
    cudaStream_t own_stream = nullptr;  // created by us
    // host-pointer calls on large batches are pipelined in slices: H2D of slice i+1 and D2H of slice i-1 overlap
    // the kernels of slice i
    static constexpr int kSlices = 16;  // most slices a call is cut into
    int slices = 8;                     // slices actually used (env GASMA_SLICES)
    bool slices_from_env = false;
    bool trace = false;                 // env GASMA_TRACE: per-call timing of the host-pointer path on stderr
    cudaEvent_t ev_t0 = nullptr, ev_t1 = nullptr, ev_t2 = nullptr;
    cudaStream_t copy_in = nullptr, copy_out = nullptr;
    cudaEvent_t ev_in[kSlices] = {}, ev_comp[kSlices] = {}, ev_off[kSlices] = {}, ev_fence = nullptr;
    int64_t launches = 0;
    std::string err;
    // device scratch
    // per-tile op counts of the greedy path, two arrays used alternately: the greedy kernel of batch i accumulates into
    // g_sums[i & 1], compact_rows_kernel of batch i reads it and zeroes the other one for batch i+1
    DevBuf g_sums[2];
    int g_dirty[2] = {0, 0};     // tiles of g_sums[q] that may be non-zero
    int g_parity = 0;            // array the NEXT greedy launch uses
    int base_parity = 0;         // misc[8 + base_parity] holds the running op total the next carry_ops step starts from
    DevBuf ops_rows;             // greedy CIGAR staging: 32-byte primary row per pair (greedy.cuh:OpSink)
    DevBuf ops_stage, nops, tile_sums, tile_bases, misc;  // misc: [0] status u32, [1] total ops u32, [2..3] step counter u64, [8..9] running op base (alternating)
    // device mirrors for the host-pointer entry points
    DevBuf d_reads, d_refs, d_read_off, d_ref_off, d_cost, d_ops, d_op_off, d_p0, d_p1;
    DevBuf gen_len, d_bound, nw_deferred, nw_tb, nw_pool_hi, nw_pool_tb;
    // gasma_benchmark_batch scratch: both CIGAR CSR arrays, penalties, counters, answers
    DevBuf bm_gops, bm_goff, bm_nops, bm_noff, bm_gcost, bm_nwpen, bm_counters, bm_answers, bm_hard, bm_nstage, bm_nnops;
    int64_t bm_gcap = 0, bm_ncap = 0;
    uint32_t* h_misc = nullptr;  // pinned, 32 words: [0..3] device misc mirror, [8..9] op totals, [10..17] counters
    // timing hooks: a ring of event pairs per dominant kernel (0 = greedy, 1 = NW, 2 = LEAP), read by gasma_kernel_time_ms
    static constexpr int kRing = 64;
    cudaEvent_t ev[3][kRing][2] = {};
    int64_t ev_count[3] = {0, 0, 0};   // launches recorded since the last read
};

namespace {

void async_destroy(gasma_ctx* ctx);   // async_entry.inl

int fail(gasma_ctx* c, int code, const char* fmt, ...) {
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    if (c) c->err = buf; else g_create_error = buf;
    return code;
}

#define CU(call)                                                                                          \
    do {                                                                                                  \
        cudaError_t _e = (call);                                                                          \
        if (_e != cudaSuccess)                                                                            \
            return fail(ctx, GASMA_E_CUDA, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(_e), __FILE__, __LINE__); \
    } while (0)

int reserve(gasma_ctx* ctx, DevBuf& b, size_t bytes) {
    if (bytes <= b.cap) return GASMA_OK;
    if (b.p) { CU(cudaStreamSynchronize(ctx->stream)); CU(cudaFree(b.p)); b.p = nullptr; b.cap = 0; }
    size_t want = bytes + bytes / 4 + 256;
    CU(cudaMalloc(&b.p, want));
    b.cap = want;
    return GASMA_OK;
}

int validate_params(const gasma_params* p, std::string& why) {
    if (!p) { why = "params is null"; return GASMA_E_ARG; }
    if (p->max_length != 64 && p->max_length != 128 && p->max_length != 192 && p->max_length != 256 &&
        p->max_length != 512 && p->max_length != 1024) {
        why = "max_length must be 64, 128, 192, 256 (one alignment per thread) or 512, 1024 (one alignment per warp)";
        return GASMA_E_UNSUPPORTED;
    }
    if (p->k < 0 || p->k > GASMA_MAX_K) { why = "k outside [0, MAX_K=50]"; return GASMA_E_BAND; }
    if (p->alignment_type != GASMA_GLOBAL && p->alignment_type != GASMA_SEMI_GLOBAL) {
        why = "alignment_type must be GLOBAL or SEMI_GLOBAL (LOCAL is unsupported in the reference too)";
        return GASMA_E_UNSUPPORTED;
    }
    if (p->x < 0 || p->o < 0 || p->e < 0 || p->x > 1000 || p->o > 1000 || p->e > 1000) { why = "penalties must be in [0,1000]"; return GASMA_E_ARG; }
    if (!(p->match_prob > 0) || !(p->mismatch_prob > 0) || !(p->indel_prob > 0)) { why = "probabilities must be > 0"; return GASMA_E_ARG; }
    return GASMA_OK;
}

void derive(gasma_ctx* c) {
    const gasma_params& p = c->params;
    c->W = p.max_length / 64;
    c->kp.k = p.k; c->kp.x = p.x; c->kp.o = p.o; c->kp.e = p.e;
    c->kp.global = p.alignment_type == GASMA_GLOBAL ? 1 : 0;
    // hurdle_matrix.h:536-538, host libm so the constants are the reference's bit for bit
    c->kp.ms = log(p.match_prob / 0.25);
    c->kp.mms = log(p.mismatch_prob / 0.25);
    c->kp.ids = log(p.indel_prob / 2 / 0.25);
    c->auto_warps = 0;
    c->kp.correction = c->correction;
    c->kp.nl = 2 * p.k + 1 + (c->correction ? c->max_delta : 0);
}

constexpr size_t kMaxSmem = 227 * 1024;

// warps per CTA of the greedy kernel: 4 when the lanes of 128 alignments fit shared memory, else 1
template <int W>
int greedy_warps_for(int nl) {
    if (greedy_smem_bytes<W>(4, nl) <= kMaxSmem) return 4;
    if (greedy_smem_bytes<W>(1, nl) <= kMaxSmem) return 1;
    return 0;
}

template <int W, int WARPS, bool CORR, int FLIP = 0>
int launch_greedy_t(gasma_ctx* ctx, const uint8_t* reads, const uint32_t* roff, const uint8_t* refs, const uint32_t* foff,
                    int64_t n, int32_t* cost, uint16_t* rows, uint16_t* stage, int stride, uint32_t* nops, uint32_t* chunk_counter,
                    uint32_t* status, unsigned long long* steps, uint32_t* tsums) {
    size_t smem = greedy_smem_bytes<W>(WARPS, ctx->kp.nl);
    if (const char* env = getenv("GASMA_SMEM_PAD")) smem += (size_t)atoi(env);   // occupancy experiments only
    auto kern = greedy_kernel<W, WARPS, CORR, FLIP>;
    CU(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int per_sm = 0;
    CU(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, 32 * WARPS, smem));
    if (per_sm < 1) return fail(ctx, GASMA_E_CUDA, "greedy kernel does not fit an SM");
    // persistent grid: every resident warp pulls 32-pair chunks until the batch is exhausted
    const int64_t chunks = (n + kPool - 1) / kPool;
    int64_t grid = (int64_t)ctx->sm_count * per_sm;
    if (grid * WARPS > chunks) grid = (chunks + WARPS - 1) / WARPS;
    if (!ctx->counter_clean) CU(cudaMemsetAsync(chunk_counter, 0, 4, ctx->stream));
    kern<<<(unsigned)grid, 32 * WARPS, smem, ctx->stream>>>(ctx->kp, reads, roff, refs, foff, (long long)n, cost, rows, stage, stride,
                                                          nops, chunk_counter, ctx->refill_threshold, status, steps, tsums,
                                                          ctx->pk_r1, ctx->pk_f1);
    CU(cudaGetLastError());
    ctx->launches++;
    return GASMA_OK;
}

template <int W>
int launch_greedy_w(gasma_ctx* ctx, int warps, const uint8_t* reads, const uint32_t* roff, const uint8_t* refs,
                    const uint32_t* foff, int64_t n, int32_t* cost, uint16_t* rows, uint16_t* stage, int stride, uint32_t* nops,
                    uint32_t* chunk_counter, uint32_t* status, unsigned long long* steps, uint32_t* tsums) {
    if (ctx->mask_variant) {   // the other mask variants: one CTA shape each, not tuned further
        if (ctx->mask_variant == 1) {
            if (warps == 4) return launch_greedy_t<W, 4, false, 1>(ctx, reads, roff, refs, foff, n, cost, rows, stage, stride, nops, chunk_counter, status, steps, tsums);
            return launch_greedy_t<W, 1, false, 1>(ctx, reads, roff, refs, foff, n, cost, rows, stage, stride, nops, chunk_counter, status, steps, tsums);
        }
        if (warps == 4) return launch_greedy_t<W, 4, false, 2>(ctx, reads, roff, refs, foff, n, cost, rows, stage, stride, nops, chunk_counter, status, steps, tsums);
        return launch_greedy_t<W, 1, false, 2>(ctx, reads, roff, refs, foff, n, cost, rows, stage, stride, nops, chunk_counter, status, steps, tsums);
    }
    if (!ctx->kp.correction && warps == 4) {
        // CTA shape: per-warp work is independent, so the shape only decides how many warps an SM holds (shared
        // memory per warp is fixed, the per-CTA reservation and the register file are not).  Pick the most resident warps.
        int pick = ctx->greedy_warps;
        if (pick == 0) {
            if (ctx->auto_warps == 0) {
                int best = 0;
                auto consider = [&](int w, auto kern) {
                    const size_t smem = greedy_smem_bytes<W>(w, ctx->kp.nl);
                    int per_sm = 0;
                    if (smem > kMaxSmem) return;
                    if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess ||
                        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, 32 * w, smem) != cudaSuccess) {
                        cudaGetLastError();
                        return;
                    }
                    if (per_sm * w > best) { best = per_sm * w; ctx->auto_warps = w; }
                };
                consider(4, greedy_kernel<W, 4, false>);
                consider(8, greedy_kernel<W, 8, false>);
                consider(2, greedy_kernel<W, 2, false>);
                if (ctx->auto_warps == 0) ctx->auto_warps = 4;
            }
            pick = ctx->auto_warps;
        }
        if (pick == 2) return launch_greedy_t<W, 2, false>(ctx, reads, roff, refs, foff, n, cost, rows, stage, stride, nops, chunk_counter, status, steps, tsums);
        if (pick == 8 && greedy_smem_bytes<W>(8, ctx->kp.nl) <= kMaxSmem)
            return launch_greedy_t<W, 8, false>(ctx, reads, roff, refs, foff, n, cost, rows, stage, stride, nops, chunk_counter, status, steps, tsums);
    }
    if (ctx->kp.correction) {
        if (warps == 4) return launch_greedy_t<W, 4, true>(ctx, reads, roff, refs, foff, n, cost, rows, stage, stride, nops, chunk_counter, status, steps, tsums);
        return launch_greedy_t<W, 1, true>(ctx, reads, roff, refs, foff, n, cost, rows, stage, stride, nops, chunk_counter, status, steps, tsums);
    }
    if (warps == 4) return launch_greedy_t<W, 4, false>(ctx, reads, roff, refs, foff, n, cost, rows, stage, stride, nops, chunk_counter, status, steps, tsums);
    return launch_greedy_t<W, 1, false>(ctx, reads, roff, refs, foff, n, cost, rows, stage, stride, nops, chunk_counter, status, steps, tsums);
}

int greedy_tile_threads(const gasma_ctx* c) {   // 0 = configuration does not fit
    switch (c->W) {
        case 1: return 32 * greedy_warps_for<1>(c->kp.nl);
        case 2: return 32 * greedy_warps_for<2>(c->kp.nl);
        case 3: return 32 * greedy_warps_for<3>(c->kp.nl);
        case 4: return 32 * greedy_warps_for<4>(c->kp.nl);
        case 8: return 4 * WarpLayout<8>::bytes_per_warp(c->kp.nl) + 64 <= kMaxSmem ? 128 : 0;
        default: return 4 * WarpLayout<16>::bytes_per_warp(c->kp.nl) + 64 <= kMaxSmem ? 128 : 0;
    }
}

// one alignment per warp (max_length 512 / 1024); ITER = lane slots per thread
template <int W, int ITER>
int launch_greedy_warp_i(gasma_ctx* ctx, const uint8_t* reads, const uint32_t* roff, const uint8_t* refs, const uint32_t* foff,
                         int64_t n, int32_t* cost, uint16_t* rows, uint16_t* stage, int stride, uint32_t* nops, uint32_t* pair_counter,
                         uint32_t* status, unsigned long long* steps, uint32_t* tsums) {
    constexpr int WARPS = 4;
    const size_t smem = WARPS * ((WarpLayout<W>::bytes_per_warp(ctx->kp.nl) + 15) / 16 * 16);
    auto kern = greedy_warp_kernel<W, WARPS, ITER>;
    CU(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int per_sm = 0;
    CU(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, 32 * WARPS, smem));
    if (per_sm < 1) return fail(ctx, GASMA_E_CUDA, "warp-per-alignment kernel does not fit an SM");
    int64_t grid = (int64_t)ctx->sm_count * per_sm;
    if (grid * WARPS > n) grid = (n + WARPS - 1) / WARPS;
    if (!ctx->counter_clean) CU(cudaMemsetAsync(pair_counter, 0, 4, ctx->stream));
    kern<<<(unsigned)grid, 32 * WARPS, smem, ctx->stream>>>(ctx->kp, reads, roff, refs, foff, (long long)n, cost, rows, stage, stride, nops,
                                                          pair_counter, status, steps, tsums);
    CU(cudaGetLastError());
    ctx->launches++;
    return GASMA_OK;
}

template <int W>
int launch_greedy_warp(gasma_ctx* ctx, const uint8_t* reads, const uint32_t* roff, const uint8_t* refs, const uint32_t* foff,
                       int64_t n, int32_t* cost, uint16_t* rows, uint16_t* stage, int stride, uint32_t* nops, uint32_t* pair_counter,
                       uint32_t* status, unsigned long long* steps, uint32_t* tsums) {
    const int nl = ctx->kp.nl;   // lane slots per alignment: 2k+1 (+ max_delta with the widened band), at most 2 * MAX_K + 1
    if (nl <= 32) return launch_greedy_warp_i<W, 1>(ctx, reads, roff, refs, foff, n, cost, rows, stage, stride, nops, pair_counter, status, steps, tsums);
    if (nl <= 64) return launch_greedy_warp_i<W, 2>(ctx, reads, roff, refs, foff, n, cost, rows, stage, stride, nops, pair_counter, status, steps, tsums);
    if (nl <= 128) return launch_greedy_warp_i<W, 4>(ctx, reads, roff, refs, foff, n, cost, rows, stage, stride, nops, pair_counter, status, steps, tsums);
    return fail(ctx, GASMA_E_BAND, "%d lane slots per alignment exceed the warp kernel's 128", nl);
}

int check_status(gasma_ctx* ctx, uint32_t st) {
    if (st & (kStatusLength | kStatusTile)) return fail(ctx, GASMA_E_LENGTH, "a string is longer than max_length=%d", ctx->params.max_length);
    if (st & kStatusBand) {
        if (ctx->correction)
            return fail(ctx, GASMA_E_BAND, "|ref_len - read_len| > max_delta=%d of the widened band, or |ref_len - read_len| + k > MAX_K=%d",
                        ctx->max_delta, GASMA_MAX_K);
        return fail(ctx, GASMA_E_BAND, "|ref_len - read_len| > MAX_K=%d (out of bounds in the reference)", GASMA_MAX_K);
    }
    if (st & kStatusSteps) return fail(ctx, GASMA_E_INTERNAL, "greedy traversal hit the step limit");
    return GASMA_OK;
}

}  // namespace

extern "C" {

void gasma_default_params(gasma_params* p) {
    if (!p) return;
    p->alignment_type = GASMA_GLOBAL;
    p->x = p->o = p->e = 1;       // benchmark.cpp:22, hurdle_matrix.h:554-556
    p->k = 3;
    p->max_length = 128;          // utils.h:23-25
    p->match_prob = 0.80;         // hurdle_matrix.h:557-559
    p->mismatch_prob = 0.20 / 3;
    p->indel_prob = 0.40 / 3;
}

int gasma_create(const gasma_params* params, int device, gasma_ctx** out) {
    gasma_ctx* ctx = nullptr;  // so CU() reports into g_create_error
    if (!out) return fail(nullptr, GASMA_E_ARG, "out is null");
    *out = nullptr;
    std::string why;
    int rc = validate_params(params, why);
    if (rc) return fail(nullptr, rc, "%s", why.c_str());
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev == 0)
        return fail(nullptr, GASMA_E_CUDA, "no CUDA device (%s); libgasma_b200 has no CPU path", cudaGetErrorString(e));
    if (device < 0 || device >= ndev) return fail(nullptr, GASMA_E_ARG, "device %d out of range (%d devices)", device, ndev);
    cudaDeviceProp prop;
    CU(cudaGetDeviceProperties(&prop, device));
    if (prop.major != 10)
        return fail(nullptr, GASMA_E_CUDA, "device %d is sm_%d%d; this library is built for sm_100a only", device, prop.major, prop.minor);
    CU(cudaSetDevice(device));
    gasma_ctx* c = new gasma_ctx();
    c->params = *params;
    c->device = device;
    c->sm_count = prop.multiProcessorCount;
    if (const char* env = getenv("GASMA_REFILL_THRESHOLD")) {   // tuning knob; results do not depend on it
        const int v = atoi(env);
        if (v >= 1 && v <= 32) c->refill_threshold = v;
    }
    if (const char* env = getenv("GASMA_GREEDY_WARPS")) {       // CTA shape of the greedy kernel; results do not depend on it
        const int v = atoi(env);
        if (v == 2 || v == 4 || v == 8) c->greedy_warps = v;
    }
    c->trace = getenv("GASMA_TRACE") != nullptr;
    if (const char* env = getenv("GASMA_SLICES")) {             // pipeline depth of the host-pointer calls; same results
        const int v = atoi(env);
        if (v >= 1 && v <= gasma_ctx::kSlices) { c->slices = v; c->slices_from_env = true; }
    }
    derive(c);
    if (greedy_tile_threads(c) == 0) {
        delete c;
        return fail(nullptr, GASMA_E_UNSUPPORTED, "k=%d with max_length=%d does not fit shared memory", params->k, params->max_length);
    }
    cudaError_t e2 = cudaStreamCreateWithFlags(&c->own_stream, cudaStreamNonBlocking);
    if (e2 != cudaSuccess) { delete c; return fail(nullptr, GASMA_E_CUDA, "cudaStreamCreate: %s", cudaGetErrorString(e2)); }
    c->stream = c->own_stream;
    bool aux_ok = cudaStreamCreateWithFlags(&c->copy_in, cudaStreamNonBlocking) == cudaSuccess &&
                  cudaStreamCreateWithFlags(&c->copy_out, cudaStreamNonBlocking) == cudaSuccess &&
                  cudaEventCreateWithFlags(&c->ev_fence, cudaEventDisableTiming) == cudaSuccess;
    for (int i = 0; i < gasma_ctx::kSlices && aux_ok; ++i)
        aux_ok = cudaEventCreateWithFlags(&c->ev_in[i], cudaEventDisableTiming) == cudaSuccess &&
                 cudaEventCreateWithFlags(&c->ev_comp[i], cudaEventDisableTiming) == cudaSuccess &&
                 cudaEventCreateWithFlags(&c->ev_off[i], cudaEventDisableTiming) == cudaSuccess;
    if (!aux_ok) { gasma_destroy(c); return fail(nullptr, GASMA_E_CUDA, "creating copy streams / events failed"); }
    if (cudaMalloc(&c->misc.p, 64) != cudaSuccess || cudaMemset(c->misc.p, 0, 64) != cudaSuccess ||
        cudaMallocHost((void**)&c->h_misc, 128) != cudaSuccess) {
        gasma_destroy(c);
        return fail(nullptr, GASMA_E_CUDA, "allocating context scratch failed");
    }
    c->misc.cap = 64;
    for (auto& kind : c->ev)
        for (auto& pair : kind)
            for (auto& ev : pair)
                if (cudaEventCreate(&ev) != cudaSuccess) { gasma_destroy(c); return fail(nullptr, GASMA_E_CUDA, "cudaEventCreate failed"); }
    *out = c;
    return GASMA_OK;
}

void gasma_destroy(gasma_ctx* c) {
    if (!c) return;
    async_destroy(c);
    cudaSetDevice(c->device);
    if (c->stream) cudaStreamSynchronize(c->stream);
    DevBuf* bufs[] = {&c->g_sums[0], &c->g_sums[1], &c->ops_rows, &c->ops_stage, &c->nops, &c->tile_sums, &c->tile_bases, &c->misc, &c->d_reads, &c->d_refs,
                      &c->d_read_off, &c->d_ref_off, &c->d_cost, &c->d_ops, &c->d_op_off, &c->d_p0, &c->d_p1, &c->gen_len, &c->d_bound, &c->nw_deferred, &c->nw_tb, &c->nw_pool_hi, &c->nw_pool_tb,
                      &c->bm_hard, &c->bm_gops, &c->bm_goff, &c->bm_nops, &c->bm_noff, &c->bm_gcost, &c->bm_nwpen, &c->bm_counters, &c->bm_answers, &c->bm_nstage, &c->bm_nnops, &c->d_r1, &c->d_f1, &c->nccl_words};
    for (DevBuf* b : bufs) if (b->p) cudaFree(b->p);
    if (c->h_misc) cudaFreeHost(c->h_misc);
    for (auto& kind : c->ev) for (auto& pair : kind) for (auto& ev : pair) if (ev) cudaEventDestroy(ev);
    for (int i = 0; i < gasma_ctx::kSlices; ++i) {
        if (c->ev_in[i]) cudaEventDestroy(c->ev_in[i]);
        if (c->ev_comp[i]) cudaEventDestroy(c->ev_comp[i]);
        if (c->ev_off[i]) cudaEventDestroy(c->ev_off[i]);
    }
    if (c->ev_fence) cudaEventDestroy(c->ev_fence);
    for (cudaEvent_t e : {c->ev_t0, c->ev_t1, c->ev_t2}) if (e) cudaEventDestroy(e);
    if (c->copy_in) cudaStreamDestroy(c->copy_in);
    if (c->copy_out) cudaStreamDestroy(c->copy_out);
    if (c->own_stream) cudaStreamDestroy(c->own_stream);
    delete c;
}

const char* gasma_last_error(const gasma_ctx* ctx) { return ctx ? ctx->err.c_str() : g_create_error.c_str(); }

int gasma_set_band(gasma_ctx* ctx, int k) {
    if (!ctx) return GASMA_E_ARG;
    if (k < 0 || k > GASMA_MAX_K) return fail(ctx, GASMA_E_BAND, "k=%d outside [0, %d]", k, GASMA_MAX_K);
    if (ctx->correction && k + ctx->max_delta > GASMA_MAX_K)
        return fail(ctx, GASMA_E_BAND, "k=%d + max_delta=%d of the widened band exceeds MAX_K=%d", k, ctx->max_delta, GASMA_MAX_K);
    int old = ctx->params.k;
    ctx->params.k = k;
    derive(ctx);
    if (greedy_tile_threads(ctx) == 0) {
        ctx->params.k = old;
        derive(ctx);
        return fail(ctx, GASMA_E_UNSUPPORTED, "k=%d with max_length=%d does not fit shared memory", k, ctx->params.max_length);
    }
    return GASMA_OK;
}

int gasma_set_correction(gasma_ctx* ctx, int enabled, int max_delta) {
    if (!ctx) return GASMA_E_ARG;
    if (enabled && (max_delta < 0 || max_delta + ctx->params.k > GASMA_MAX_K))
        return fail(ctx, GASMA_E_BAND, "max_delta=%d: need 0 <= max_delta and max_delta + k <= MAX_K=%d", max_delta, GASMA_MAX_K);
    if (enabled && ctx->mask_variant) return fail(ctx, GASMA_E_UNSUPPORTED, "the widened band is not combined with the other mask variants");
    const int old_c = ctx->correction, old_d = ctx->max_delta;
    ctx->correction = enabled ? 1 : 0;
    ctx->max_delta = enabled ? max_delta : 0;
    derive(ctx);
    if (greedy_tile_threads(ctx) == 0) {
        ctx->correction = old_c; ctx->max_delta = old_d;
        derive(ctx);
        return fail(ctx, GASMA_E_UNSUPPORTED, "k=%d + max_delta=%d lanes with max_length=%d do not fit shared memory", ctx->params.k,
                    max_delta, ctx->params.max_length);
    }
    return GASMA_OK;
}

int gasma_set_mask_variant(gasma_ctx* ctx, int variant) {
    if (!ctx) return GASMA_E_ARG;
    if (variant < 0 || variant > 2) return fail(ctx, GASMA_E_ARG, "mask variant must be 0 (flip_short_hurdles(1)), 1 (flip_short_hurdles(2)) or 2 (+ flip_short_matches(1))");
    if (variant && ctx->W > 4) return fail(ctx, GASMA_E_UNSUPPORTED, "the other mask variants cover max_length <= 256");
    if (variant && ctx->correction) return fail(ctx, GASMA_E_UNSUPPORTED, "the other mask variants are not combined with the widened band");
    ctx->mask_variant = variant;
    return GASMA_OK;
}

int gasma_set_stream(gasma_ctx* ctx, void* cuda_stream) {
    if (!ctx) return GASMA_E_ARG;
    CU(cudaStreamSynchronize(ctx->stream));
    ctx->stream = (cudaStream_t)cuda_stream;  // taken literally: NULL is the legacy default stream
    return GASMA_OK;
}

int gasma_synchronize(gasma_ctx* ctx) {
    if (!ctx) return GASMA_E_ARG;
    CU(cudaSetDevice(ctx->device));
    CU(cudaMemcpyAsync(ctx->h_misc, ctx->misc.p, 16, cudaMemcpyDeviceToHost, ctx->stream));
    CU(cudaMemsetAsync(ctx->misc.p, 0, 4, ctx->stream));  // status is consumed by this call
    CU(cudaStreamSynchronize(ctx->stream));
    return check_status(ctx, ctx->h_misc[0]);
}

int64_t gasma_launch_count(const gasma_ctx* ctx) { return ctx ? ctx->launches : 0; }

double gasma_kernel_time_ms(gasma_ctx* ctx, int which, int64_t* launches) {
    if (launches) *launches = 0;
    if (!ctx || which < 0 || which > 2) return -1.0;
    const int64_t cnt = ctx->ev_count[which];
    const int64_t take = cnt < gasma_ctx::kRing ? cnt : gasma_ctx::kRing;
    double total = 0.0;
    for (int64_t i = cnt - take; i < cnt; ++i) {
        cudaEvent_t* e = ctx->ev[which][i % gasma_ctx::kRing];
        float ms = 0.f;
        if (cudaEventSynchronize(e[1]) != cudaSuccess || cudaEventElapsedTime(&ms, e[0], e[1]) != cudaSuccess) return -1.0;
        total += ms;
    }
    ctx->ev_count[which] = 0;
    if (launches) *launches = take;
    return total;
}

int64_t gasma_last_step_count(gasma_ctx* ctx) {
    if (!ctx) return -1;
    unsigned long long v = 0;
    if (cudaMemcpyAsync(&v, (uint32_t*)ctx->misc.p + 2, 8, cudaMemcpyDeviceToHost, ctx->stream) != cudaSuccess) return -1;
    if (cudaMemsetAsync((uint32_t*)ctx->misc.p + 2, 0, 8, ctx->stream) != cudaSuccess) return -1;
    if (cudaStreamSynchronize(ctx->stream) != cudaSuccess) return -1;
    return (int64_t)v;
}

int gasma_render_cigar(const gasma_cigar_op* ops, int n_ops, char* out, int cap) {
    static const char letters[] = "MID=X???";
    int pos = 0;
    for (int i = 0; i < n_ops; ++i) {
        char tmp[16];
        int c = snprintf(tmp, sizeof tmp, "%d%c", (int)GASMA_OP_LEN(ops[i]), letters[GASMA_OP_CODE(ops[i])]);
        for (int j = 0; j < c; ++j)
            if (out && pos + j < cap - 1) out[pos + j] = tmp[j];
        pos += c;
    }
    if (out && cap > 0) out[pos < cap - 1 ? pos : cap - 1] = '\0';
    return pos;
}

}  // extern "C"

// =================================================================================================
// K1 / K2+K3 entry points
// =================================================================================================
namespace {

bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }

template <int W>
int launch_pack_w(gasma_ctx* ctx, const uint8_t* seqs, const uint32_t* off, int64_t n, uint64_t* p0, uint64_t* p1) {
    const unsigned grid = (unsigned)((n + kTile - 1) / kTile);
    pack2bit_kernel<W, kTile><<<grid, kTile, 0, ctx->stream>>>(seqs, off, (long long)n, p0, p1, (uint32_t*)ctx->misc.p);
    CU(cudaGetLastError());
    ctx->launches++;
    return GASMA_OK;
}

// How the CIGAR ops leave a greedy step:
//   d_ops == nullptr, rows_only == false   costs only
//   d_ops != nullptr                       CSR: compact_rows_kernel writes d_ops / d_op_off
//   rows_only                              left in the staging rows + ctx->nops (the benchmark body's verifier reads them there)
// carry_ops: op_off continues from the running total left by the previous call (slices of one batch share one
// CSR array); otherwise it starts at 0.
constexpr int64_t kMaxCsrPairs = (int64_t)4096 * kCsrTile;   // pairs per compact_rows_kernel launch (a CTA adds up <= 4096 tile counts)

int greedy_dev_impl(gasma_ctx* ctx, const char* d_reads, const uint32_t* d_read_off, const char* d_refs,
                    const uint32_t* d_ref_off, int64_t n, int32_t* d_cost, gasma_cigar_op* d_ops, int64_t ops_cap,
                    uint32_t* d_op_off, bool carry_ops = false, bool rows_only = false) {
    if (!ctx) return GASMA_E_ARG;
    if (n < 0 || n > (int64_t)1 << 31) return fail(ctx, GASMA_E_ARG, "n_pairs out of range");
    if (!d_read_off || !d_ref_off || !d_cost) return fail(ctx, GASMA_E_ARG, "null offsets/cost pointer");
    if ((d_ops != nullptr) != (d_op_off != nullptr)) return fail(ctx, GASMA_E_ARG, "ops and op_off must be given together");
    if (!aligned16(d_reads) || !aligned16(d_refs)) return fail(ctx, GASMA_E_ARG, "sequence buffers must be 16-byte aligned");
    CU(cudaSetDevice(ctx->device));
    if (n == 0) {
        if (d_op_off && !carry_ops) CU(cudaMemsetAsync(d_op_off, 0, 4, ctx->stream));
        return GASMA_OK;
    }
    const bool want_csr = d_ops != nullptr;
    if (want_csr && n > kMaxCsrPairs) {   // very large batches: several steps that share the CSR array
        for (int64_t lo = 0; lo < n; lo += kMaxCsrPairs) {
            const int64_t cnt = n - lo < kMaxCsrPairs ? n - lo : kMaxCsrPairs;
            const int rc = greedy_dev_impl(ctx, d_reads, d_read_off + lo, d_refs, d_ref_off + lo, cnt, d_cost + lo, d_ops, ops_cap,
                                           d_op_off + lo, carry_ops || lo > 0, false);
            if (rc) return rc;
        }
        return GASMA_OK;
    }
    if (ctx->pk_r1 && ctx->W > 4)
        return fail(ctx, GASMA_E_UNSUPPORTED, "packed input covers max_length <= 256 (one alignment per thread)");
    const int warps = greedy_tile_threads(ctx) / 32;
    const int64_t ntiles = (n + kCsrTile - 1) / kCsrTile;
    const int stride = 2 * ctx->params.max_length + 8;
    uint32_t* misc = (uint32_t*)ctx->misc.p;
    const bool want_ops = want_csr || rows_only;
    const int q = ctx->g_parity;
    int rc;
    if ((rc = reserve(ctx, ctx->ops_rows, (size_t)n * kRowPitch * sizeof(uint16_t)))) return rc;   // always: emit() is unconditional
    if (want_ops) {
        if ((rc = reserve(ctx, ctx->ops_stage, (size_t)n * stride * sizeof(uint16_t)))) return rc;
        if ((rc = reserve(ctx, ctx->nops, (size_t)n * 4))) return rc;
    }
    if (want_csr) {
        for (int j = 0; j < 2; ++j) {
            const void* before = ctx->g_sums[j].p;
            if ((rc = reserve(ctx, ctx->g_sums[j], (size_t)ntiles * 4))) return rc;
            if (ctx->g_sums[j].p != before) {
                CU(cudaMemsetAsync(ctx->g_sums[j].p, 0, ctx->g_sums[j].cap, ctx->stream));
                ctx->g_dirty[j] = 0;
            }
        }
    }
    uint32_t* tsums = want_csr ? (uint32_t*)ctx->g_sums[q].p : nullptr;
    uint16_t* rows = (uint16_t*)ctx->ops_rows.p;
    uint16_t* stage = want_ops ? (uint16_t*)ctx->ops_stage.p : nullptr;
    uint32_t* nops = want_ops ? (uint32_t*)ctx->nops.p : nullptr;
    const uint8_t* r = (const uint8_t*)d_reads;
    const uint8_t* f = (const uint8_t*)d_refs;
    unsigned long long* steps = (unsigned long long*)(misc + 2);
    uint32_t* chunk_counter = misc + 6;
    cudaEvent_t* evp = ctx->ev[0][ctx->ev_count[0] % gasma_ctx::kRing];
    CU(cudaEventRecord(evp[0], ctx->stream));
    switch (ctx->W) {
        case 1: rc = launch_greedy_w<1>(ctx, warps, r, d_read_off, f, d_ref_off, n, d_cost, rows, stage, stride, nops, chunk_counter, misc, steps, tsums); break;
        case 2: rc = launch_greedy_w<2>(ctx, warps, r, d_read_off, f, d_ref_off, n, d_cost, rows, stage, stride, nops, chunk_counter, misc, steps, tsums); break;
        case 3: rc = launch_greedy_w<3>(ctx, warps, r, d_read_off, f, d_ref_off, n, d_cost, rows, stage, stride, nops, chunk_counter, misc, steps, tsums); break;
        case 4: rc = launch_greedy_w<4>(ctx, warps, r, d_read_off, f, d_ref_off, n, d_cost, rows, stage, stride, nops, chunk_counter, misc, steps, tsums); break;
        case 8: rc = launch_greedy_warp<8>(ctx, r, d_read_off, f, d_ref_off, n, d_cost, rows, stage, stride, nops, chunk_counter, misc, steps, tsums); break;
        default: rc = launch_greedy_warp<16>(ctx, r, d_read_off, f, d_ref_off, n, d_cost, rows, stage, stride, nops, chunk_counter, misc, steps, tsums); break;
    }
    if (rc) return rc;
    CU(cudaEventRecord(evp[1], ctx->stream));
    ctx->ev_count[0]++;
    ctx->counter_clean = false;   // the kernel leaves its chunk counter past the end
    if (want_csr) {
        if (ntiles > ctx->g_dirty[q]) ctx->g_dirty[q] = (int)ntiles;
        // one kernel: tile bases from the counts the greedy kernel accumulated, CSR offsets, coalesced op copy; it also
        // zeroes the other count array for the next batch and resets the chunk counter
        const int bp = ctx->base_parity;
        compact_rows_kernel<kCsrTile><<<(unsigned)ntiles, kCsrTile, 0, ctx->stream>>>(
            rows, stage, stride, nops, tsums, (uint32_t*)ctx->g_sums[q ^ 1].p, ctx->g_dirty[q ^ 1],
            carry_ops ? misc + 8 + bp : nullptr, misc + 8 + (bp ^ 1), (long long)n, d_ops, (long long)ops_cap, d_op_off, chunk_counter);
        CU(cudaGetLastError());
        ctx->g_dirty[q ^ 1] = 0;
        ctx->g_parity = q ^ 1;
        ctx->base_parity = bp ^ 1;
        ctx->counter_clean = true;
        ctx->launches += 1;
    }
    return GASMA_OK;
}

int check_offsets(gasma_ctx* ctx, const uint32_t* off, int64_t n, const char* what);

// Host-pointer greedy call on a large batch, pipelined over kSlices slices of pairs (slice edges on multiples of
// 128 pairs).  Device mirrors are already reserved by the caller.  Streams: copy_in (H2D), ctx->stream (kernels),
// copy_out (D2H); the CSR op array is shared by all slices, op_off continues across them on the device.
int greedy_batch_pipelined_body(gasma_ctx* ctx, const char* reads, const uint32_t* read_off, const char* refs,
                                const uint32_t* ref_off, int64_t n, int32_t* cost, gasma_cigar_op* ops, int64_t ops_cap,
                                uint32_t* op_off, const uint16_t* reads_p1, const uint16_t* refs_p1);

// Every error exit of the pipelined call goes through here: copies from / into the caller's buffers may still be
// queued on the copy streams, and the per-tile op counts may be half accumulated.  Drain all three streams and put
// the context back into its clean state before the caller gets its buffers back.
// reads_p1 / refs_p1 != nullptr: packed input -- reads / refs are the HOST plane-0 streams (uint16 per 16 characters),
// these the plane-1 streams (gasma_greedy_batch_packed); the offsets are the same CSR offsets in characters.
int greedy_batch_pipelined(gasma_ctx* ctx, const char* reads, const uint32_t* read_off, const char* refs,
                           const uint32_t* ref_off, int64_t n, int32_t* cost, gasma_cigar_op* ops, int64_t ops_cap,
                           uint32_t* op_off, const uint16_t* reads_p1 = nullptr, const uint16_t* refs_p1 = nullptr) {
    const int rc = greedy_batch_pipelined_body(ctx, reads, read_off, refs, ref_off, n, cost, ops, ops_cap, op_off, reads_p1, refs_p1);
    ctx->pk_r1 = ctx->pk_f1 = nullptr;
    if (rc != GASMA_OK && rc != GASMA_E_CAPACITY) {
        const std::string keep = ctx->err;
        cudaStreamSynchronize(ctx->copy_in);
        cudaStreamSynchronize(ctx->stream);
        cudaStreamSynchronize(ctx->copy_out);
        for (int j = 0; j < 2; ++j) {
            if (ctx->g_sums[j].p) cudaMemsetAsync(ctx->g_sums[j].p, 0, ctx->g_sums[j].cap, ctx->stream);
            ctx->g_dirty[j] = 0;
        }
        cudaMemsetAsync(ctx->misc.p, 0, 4, ctx->stream);   // status word
        cudaStreamSynchronize(ctx->stream);
        cudaGetLastError();
        ctx->counter_clean = false;
        ctx->err = keep;
    }
    return rc;
}

int greedy_batch_pipelined_body(gasma_ctx* ctx, const char* reads, const uint32_t* read_off, const char* refs,
                                const uint32_t* ref_off, int64_t n, int32_t* cost, gasma_cigar_op* ops, int64_t ops_cap,
                                uint32_t* op_off, const uint16_t* reads_p1, const uint16_t* refs_p1) {
    const bool packed = reads_p1 != nullptr;
    // packed input is 4x smaller and comes as four arrays: with 8 slices its copies are ~2.5 MB each and the per-copy
    // overhead shows (measured per 1 M pairs: 4 slices 2.04 ms, 8: 2.21, 12: 2.52, 16: 2.96); GASMA_SLICES overrides
    const int S = packed && !ctx->slices_from_env ? 4 : ctx->slices;
    const size_t ob = (size_t)(n + 1) * 4;
    int64_t edge[gasma_ctx::kSlices + 1];
    // The call ends one slice's kernels + result copies after the last input byte has arrived, so the last two slices
    // are short (weights 1, ..., 1, 1/2, 1/5): the tail shrinks with them while the link stays saturated.
    {
        double w[gasma_ctx::kSlices], total = 0.0, acc = 0.0;
        for (int i = 0; i < S; ++i) { w[i] = S >= 4 && i == S - 1 ? 0.2 : (S >= 4 && i == S - 2 ? 0.5 : 1.0); total += w[i]; }
        edge[0] = 0;
        for (int i = 1; i <= S; ++i) {
            acc += w[i - 1];
            edge[i] = i == S ? n : ((int64_t)((double)n * acc / total) / 128) * 128;
            if (edge[i] < edge[i - 1]) edge[i] = edge[i - 1];
        }
    }
    char* d_reads = (char*)ctx->d_reads.p;
    char* d_refs = (char*)ctx->d_refs.p;
    uint32_t* d_roff = (uint32_t*)ctx->d_read_off.p;
    uint32_t* d_foff = (uint32_t*)ctx->d_ref_off.p;
    int32_t* d_cost = (int32_t*)ctx->d_cost.p;
    gasma_cigar_op* d_ops = ops ? (gasma_cigar_op*)ctx->d_ops.p : nullptr;
    uint32_t* d_ooff = ops ? (uint32_t*)ctx->d_op_off.p : nullptr;
    uint32_t* misc = (uint32_t*)ctx->misc.p;
    // the mirrors may still be in use by earlier work on the compute stream
    CU(cudaEventRecord(ctx->ev_fence, ctx->stream));
    CU(cudaStreamWaitEvent(ctx->copy_in, ctx->ev_fence, 0));
    CU(cudaStreamWaitEvent(ctx->copy_out, ctx->ev_fence, 0));
    const auto wall0 = std::chrono::steady_clock::now();
    if (ctx->trace) {
        if (!ctx->ev_t0) { CU(cudaEventCreate(&ctx->ev_t0)); CU(cudaEventCreate(&ctx->ev_t1)); CU(cudaEventCreate(&ctx->ev_t2)); }
        CU(cudaEventRecord(ctx->ev_t0, ctx->copy_in));
    }
    CU(cudaMemcpyAsync(d_roff, read_off, ob, cudaMemcpyHostToDevice, ctx->copy_in));
    CU(cudaMemcpyAsync(d_foff, ref_off, ob, cudaMemcpyHostToDevice, ctx->copy_in));
    for (int i = 0; i < S; ++i)   // the copies below only trust the slice edges; everything else is validated while they run
        if (read_off[edge[i]] > read_off[edge[i + 1]] || ref_off[edge[i]] > ref_off[edge[i + 1]])
            return fail(ctx, GASMA_E_ARG, "offsets are not non-decreasing around pair %lld", (long long)edge[i]);
    for (int i = 0; i < S; ++i) {
        const size_t r0 = read_off[edge[i]], r1 = read_off[edge[i + 1]], f0 = ref_off[edge[i]], f1 = ref_off[edge[i + 1]];
        if (!packed) {
            if (r1 > r0) CU(cudaMemcpyAsync(d_reads + r0, reads + r0, r1 - r0, cudaMemcpyHostToDevice, ctx->copy_in));
            if (f1 > f0) CU(cudaMemcpyAsync(d_refs + f0, refs + f0, f1 - f0, cudaMemcpyHostToDevice, ctx->copy_in));
        } else {
            // the slice's chunk range of all four streams (2 bytes per 16 characters); a chunk shared with the
            // neighbouring slice is simply copied by both
            const size_t rc0 = 2 * (r0 >> 4), rc1 = 2 * ((r1 + 15) >> 4), fc0 = 2 * (f0 >> 4), fc1 = 2 * ((f1 + 15) >> 4);
            if (rc1 > rc0) {
                CU(cudaMemcpyAsync(d_reads + rc0, reads + rc0, rc1 - rc0, cudaMemcpyHostToDevice, ctx->copy_in));
                CU(cudaMemcpyAsync((char*)ctx->d_r1.p + rc0, (const char*)reads_p1 + rc0, rc1 - rc0, cudaMemcpyHostToDevice, ctx->copy_in));
            }
            if (fc1 > fc0) {
                CU(cudaMemcpyAsync(d_refs + fc0, refs + fc0, fc1 - fc0, cudaMemcpyHostToDevice, ctx->copy_in));
                CU(cudaMemcpyAsync((char*)ctx->d_f1.p + fc0, (const char*)refs_p1 + fc0, fc1 - fc0, cudaMemcpyHostToDevice, ctx->copy_in));
            }
        }
        CU(cudaEventRecord(ctx->ev_in[i], ctx->copy_in));
    }
    if (ctx->trace) CU(cudaEventRecord(ctx->ev_t1, ctx->copy_in));
    {   // full validation of both offset arrays, hidden behind the first slice's DMA (8 MB of host reads for 1 M pairs)
        int vrc = check_offsets(ctx, read_off, n, "read");
        if (!vrc) vrc = check_offsets(ctx, ref_off, n, "ref");
        if (vrc) return vrc;
    }
    CU(cudaMemsetAsync(misc + 8 + ctx->base_parity, 0, 4, ctx->stream));   // running op total the first slice starts from
    if (packed) { ctx->pk_r1 = (const uint16_t*)ctx->d_r1.p; ctx->pk_f1 = (const uint16_t*)ctx->d_f1.p; }
    for (int i = 0; i < S; ++i) {
        const int64_t lo = edge[i], cnt = edge[i + 1] - edge[i];
        if (cnt == 0) { CU(cudaEventRecord(ctx->ev_off[i], ctx->copy_out)); continue; }
        CU(cudaStreamWaitEvent(ctx->stream, ctx->ev_in[i], 0));
        int rc = greedy_dev_impl(ctx, d_reads, d_roff + lo, d_refs, d_foff + lo, cnt, d_cost + lo, d_ops, ops_cap,
                                 d_ooff ? d_ooff + lo : nullptr, /*carry_ops=*/true);
        if (rc) return rc;
        CU(cudaEventRecord(ctx->ev_comp[i], ctx->stream));
        CU(cudaStreamWaitEvent(ctx->copy_out, ctx->ev_comp[i], 0));
        CU(cudaMemcpyAsync(cost + lo, d_cost + lo, (size_t)cnt * 4, cudaMemcpyDeviceToHost, ctx->copy_out));
        if (ops) CU(cudaMemcpyAsync(op_off + lo, d_ooff + lo, (size_t)(cnt + 1) * 4, cudaMemcpyDeviceToHost, ctx->copy_out));
        CU(cudaEventRecord(ctx->ev_off[i], ctx->copy_out));
    }
    bool overflow = false;
    if (ops) {
        for (int i = 0; i < S; ++i) {   // as soon as a slice's offsets are on the host its ops can follow
            if (edge[i + 1] == edge[i]) continue;
            CU(cudaEventSynchronize(ctx->ev_off[i]));
            const int64_t lo = op_off[edge[i]], hi = op_off[edge[i + 1]];
            if (hi > ops_cap) { overflow = true; continue; }
            if (hi > lo) CU(cudaMemcpyAsync(ops + lo, d_ops + lo, (size_t)(hi - lo) * 2, cudaMemcpyDeviceToHost, ctx->copy_out));
        }
    }
    CU(cudaStreamSynchronize(ctx->copy_out));
    int rc = gasma_synchronize(ctx);
    if (rc) return rc;
    if (ctx->trace) {
        float h2d_ms = 0.f, all_ms = 0.f;
        cudaEventRecord(ctx->ev_t2, ctx->copy_out);
        cudaEventSynchronize(ctx->ev_t2);
        cudaEventElapsedTime(&h2d_ms, ctx->ev_t0, ctx->ev_t1);
        cudaEventElapsedTime(&all_ms, ctx->ev_t0, ctx->ev_t2);
        const double wall = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - wall0).count();
        fprintf(stderr, "[gasma] pipelined call: %d slices, H2D %.3f ms (%.1f GB/s), first copy -> last offsets on host %.3f ms, wall %.3f ms\n",
                S, h2d_ms, (double)((packed ? (read_off[n] + ref_off[n]) / 4 : read_off[n] + ref_off[n]) + 2 * ob) / h2d_ms / 1e6, all_ms, wall);
    }
    if (overflow)
        return fail(ctx, GASMA_E_CAPACITY, "ops buffer holds %lld ops, %lld needed", (long long)ops_cap, (long long)op_off[n]);
    return GASMA_OK;
}

// host-side validation of a CSR offset array.  This runs before the first byte is copied, so it must not cost
// milliseconds on a 1 M-pair batch: blocks are checked with a branch-free reduction the compiler vectorises, and
// only a block that failed is searched for the position.
int check_offsets(gasma_ctx* ctx, const uint32_t* off, int64_t n, const char* what) {
    constexpr int64_t kBlock = 4096;
    for (int64_t b = 0; b < n; b += kBlock) {
        const int64_t e = b + kBlock < n ? b + kBlock : n;
        uint32_t bad = 0;
        for (int64_t i = b; i < e; ++i) bad |= (uint32_t)(off[i + 1] < off[i]);
        if (bad)
            for (int64_t i = b; i < e; ++i)
                if (off[i + 1] < off[i])
                    return fail(ctx, GASMA_E_ARG, "%s offsets are not non-decreasing at %lld", what, (long long)i);
    }
    return GASMA_OK;
}

}  // namespace

extern "C" {

int gasma_pack2bit_batch_dev(gasma_ctx* ctx, const char* d_seqs, const uint32_t* d_off, int64_t n,
                             uint64_t* d_plane0, uint64_t* d_plane1) {
    if (!ctx) return GASMA_E_ARG;
    if (n < 0 || !d_off || !d_plane0 || !d_plane1) return fail(ctx, GASMA_E_ARG, "bad argument");
    if (!aligned16(d_seqs)) return fail(ctx, GASMA_E_ARG, "sequence buffer must be 16-byte aligned");
    CU(cudaSetDevice(ctx->device));
    if (n == 0) return GASMA_OK;
    const uint8_t* s = (const uint8_t*)d_seqs;
    switch (ctx->W) {
        case 1: return launch_pack_w<1>(ctx, s, d_off, n, d_plane0, d_plane1);
        case 2: return launch_pack_w<2>(ctx, s, d_off, n, d_plane0, d_plane1);
        case 3: return launch_pack_w<3>(ctx, s, d_off, n, d_plane0, d_plane1);
        case 4: return launch_pack_w<4>(ctx, s, d_off, n, d_plane0, d_plane1);
        case 8: return launch_pack_w<8>(ctx, s, d_off, n, d_plane0, d_plane1);
        default: return launch_pack_w<16>(ctx, s, d_off, n, d_plane0, d_plane1);
    }
}

int gasma_pack2bit_batch(gasma_ctx* ctx, const char* seqs, const uint32_t* off, int64_t n, uint64_t* plane0,
                         uint64_t* plane1) {
    if (!ctx) return GASMA_E_ARG;
    if (n < 0 || !off || !plane0 || !plane1) return fail(ctx, GASMA_E_ARG, "bad argument");
    if (n == 0) return GASMA_OK;
    int rc = check_offsets(ctx, off, n, "seq");
    if (rc) return rc;
    CU(cudaSetDevice(ctx->device));
    const size_t bytes = off[n], pb = (size_t)n * ctx->W * 8;
    if ((rc = reserve(ctx, ctx->d_reads, bytes + 16))) return rc;
    if ((rc = reserve(ctx, ctx->d_read_off, (size_t)(n + 1) * 4))) return rc;
    if ((rc = reserve(ctx, ctx->d_p0, pb))) return rc;
    if ((rc = reserve(ctx, ctx->d_p1, pb))) return rc;
    if (bytes) CU(cudaMemcpyAsync(ctx->d_reads.p, seqs, bytes, cudaMemcpyHostToDevice, ctx->stream));
    CU(cudaMemcpyAsync(ctx->d_read_off.p, off, (size_t)(n + 1) * 4, cudaMemcpyHostToDevice, ctx->stream));
    rc = gasma_pack2bit_batch_dev(ctx, (const char*)ctx->d_reads.p, (const uint32_t*)ctx->d_read_off.p, n,
                                  (uint64_t*)ctx->d_p0.p, (uint64_t*)ctx->d_p1.p);
    if (rc) return rc;
    CU(cudaMemcpyAsync(plane0, ctx->d_p0.p, pb, cudaMemcpyDeviceToHost, ctx->stream));
    CU(cudaMemcpyAsync(plane1, ctx->d_p1.p, pb, cudaMemcpyDeviceToHost, ctx->stream));
    return gasma_synchronize(ctx);
}

int gasma_greedy_batch_dev(gasma_ctx* ctx, const char* d_reads, const uint32_t* d_read_off, const char* d_refs,
                           const uint32_t* d_ref_off, int64_t n_pairs, int32_t* d_cost, gasma_cigar_op* d_ops,
                           int64_t ops_cap, uint32_t* d_op_off) {
    return greedy_dev_impl(ctx, d_reads, d_read_off, d_refs, d_ref_off, n_pairs, d_cost, d_ops, ops_cap, d_op_off);
}

int gasma_greedy_batch(gasma_ctx* ctx, const char* reads, const uint32_t* read_off, const char* refs,
                       const uint32_t* ref_off, int64_t n, int32_t* cost, gasma_cigar_op* ops, int64_t ops_cap,
                       uint32_t* op_off) {
    if (!ctx) return GASMA_E_ARG;
    if (n < 0 || !read_off || !ref_off || !cost) return fail(ctx, GASMA_E_ARG, "bad argument");
    if ((ops != nullptr) != (op_off != nullptr)) return fail(ctx, GASMA_E_ARG, "ops and op_off must be given together");
    if (n == 0) { if (op_off) op_off[0] = 0; return GASMA_OK; }
    int rc;
    const bool pipelined = n >= 65536;
    if (!pipelined) {   // the pipelined path validates while its first copies are in flight
        if ((rc = check_offsets(ctx, read_off, n, "read"))) return rc;
        if ((rc = check_offsets(ctx, ref_off, n, "ref"))) return rc;
    }
    CU(cudaSetDevice(ctx->device));
    const size_t rb = read_off[n], fb = ref_off[n], ob = (size_t)(n + 1) * 4;
    if ((rc = reserve(ctx, ctx->d_reads, rb + 16))) return rc;
    if ((rc = reserve(ctx, ctx->d_refs, fb + 16))) return rc;
    if ((rc = reserve(ctx, ctx->d_read_off, ob))) return rc;
    if ((rc = reserve(ctx, ctx->d_ref_off, ob))) return rc;
    if ((rc = reserve(ctx, ctx->d_cost, (size_t)n * 4))) return rc;
    if (ops) {
        if ((rc = reserve(ctx, ctx->d_op_off, ob))) return rc;
        if ((rc = reserve(ctx, ctx->d_ops, (size_t)(ops_cap > 0 ? ops_cap : 1) * 2))) return rc;
    }
    if (pipelined) return greedy_batch_pipelined(ctx, reads, read_off, refs, ref_off, n, cost, ops, ops_cap, op_off);
    if (rb) CU(cudaMemcpyAsync(ctx->d_reads.p, reads, rb, cudaMemcpyHostToDevice, ctx->stream));
    if (fb) CU(cudaMemcpyAsync(ctx->d_refs.p, refs, fb, cudaMemcpyHostToDevice, ctx->stream));
    CU(cudaMemcpyAsync(ctx->d_read_off.p, read_off, ob, cudaMemcpyHostToDevice, ctx->stream));
    CU(cudaMemcpyAsync(ctx->d_ref_off.p, ref_off, ob, cudaMemcpyHostToDevice, ctx->stream));
    rc = greedy_dev_impl(ctx, (const char*)ctx->d_reads.p, (const uint32_t*)ctx->d_read_off.p, (const char*)ctx->d_refs.p,
                         (const uint32_t*)ctx->d_ref_off.p, n, (int32_t*)ctx->d_cost.p,
                         ops ? (gasma_cigar_op*)ctx->d_ops.p : nullptr, ops_cap, ops ? (uint32_t*)ctx->d_op_off.p : nullptr);
    if (rc) return rc;
    CU(cudaMemcpyAsync(cost, ctx->d_cost.p, (size_t)n * 4, cudaMemcpyDeviceToHost, ctx->stream));
    if (ops) CU(cudaMemcpyAsync(op_off, ctx->d_op_off.p, ob, cudaMemcpyDeviceToHost, ctx->stream));
    if ((rc = gasma_synchronize(ctx))) return rc;
    if (ops) {
        const int64_t total = op_off[n];
        if (total > ops_cap) return fail(ctx, GASMA_E_CAPACITY, "ops buffer holds %lld ops, %lld needed", (long long)ops_cap, (long long)total);
        if (total) CU(cudaMemcpyAsync(ops, ctx->d_ops.p, (size_t)total * 2, cudaMemcpyDeviceToHost, ctx->stream));
        CU(cudaStreamSynchronize(ctx->stream));
    }
    return GASMA_OK;
}

// ---- packed input: 2-bit streams ---------------------------------------------------------------
int gasma_pack2bit_stream_dev(gasma_ctx* ctx, const char* d_seqs, int64_t total_chars, uint16_t* d_plane0, uint16_t* d_plane1) {
    if (!ctx) return GASMA_E_ARG;
    if (total_chars < 0 || total_chars >= ((int64_t)1 << 32) || !d_plane0 || !d_plane1) return fail(ctx, GASMA_E_ARG, "bad argument");
    if (!aligned16(d_seqs)) return fail(ctx, GASMA_E_ARG, "sequence buffer must be 16-byte aligned");
    CU(cudaSetDevice(ctx->device));
    if (total_chars == 0) return GASMA_OK;
    const int64_t nchunks = (total_chars + 15) / 16;
    int64_t grid = (nchunks + 255) / 256;
    if (grid > (int64_t)ctx->sm_count * 16) grid = (int64_t)ctx->sm_count * 16;
    pack2bit_stream_kernel<<<(unsigned)grid, 256, 0, ctx->stream>>>((const uint8_t*)d_seqs, (uint32_t)total_chars, d_plane0, d_plane1);
    CU(cudaGetLastError());
    ctx->launches++;
    return GASMA_OK;
}

int gasma_pack2bit_stream(gasma_ctx* ctx, const char* seqs, int64_t total_chars, uint16_t* plane0, uint16_t* plane1) {
    if (!ctx) return GASMA_E_ARG;
    if (total_chars < 0 || total_chars >= ((int64_t)1 << 32) || !plane0 || !plane1 || (total_chars && !seqs)) return fail(ctx, GASMA_E_ARG, "bad argument");
    if (total_chars == 0) return GASMA_OK;
    CU(cudaSetDevice(ctx->device));
    const size_t sb = (size_t)((total_chars + 15) / 16) * 2;
    int rc;
    if ((rc = reserve(ctx, ctx->d_reads, (size_t)total_chars + 16))) return rc;
    if ((rc = reserve(ctx, ctx->d_p0, sb))) return rc;
    if ((rc = reserve(ctx, ctx->d_p1, sb))) return rc;
    CU(cudaMemcpyAsync(ctx->d_reads.p, seqs, (size_t)total_chars, cudaMemcpyHostToDevice, ctx->stream));
    if ((rc = gasma_pack2bit_stream_dev(ctx, (const char*)ctx->d_reads.p, total_chars, (uint16_t*)ctx->d_p0.p, (uint16_t*)ctx->d_p1.p))) return rc;
    CU(cudaMemcpyAsync(plane0, ctx->d_p0.p, sb, cudaMemcpyDeviceToHost, ctx->stream));
    CU(cudaMemcpyAsync(plane1, ctx->d_p1.p, sb, cudaMemcpyDeviceToHost, ctx->stream));
    return gasma_synchronize(ctx);
}

int gasma_greedy_batch_packed_dev(gasma_ctx* ctx, const uint16_t* d_read_p0, const uint16_t* d_read_p1, const uint32_t* d_read_off,
                                  const uint16_t* d_ref_p0, const uint16_t* d_ref_p1, const uint32_t* d_ref_off, int64_t n_pairs,
                                  int32_t* d_cost, gasma_cigar_op* d_ops, int64_t ops_cap, uint32_t* d_op_off) {
    if (!ctx) return GASMA_E_ARG;
    if (!d_read_p0 || !d_read_p1 || !d_ref_p0 || !d_ref_p1) return fail(ctx, GASMA_E_ARG, "null stream pointer");
    ctx->pk_r1 = d_read_p1;
    ctx->pk_f1 = d_ref_p1;
    const int rc = greedy_dev_impl(ctx, (const char*)d_read_p0, d_read_off, (const char*)d_ref_p0, d_ref_off, n_pairs, d_cost, d_ops, ops_cap, d_op_off);
    ctx->pk_r1 = ctx->pk_f1 = nullptr;
    return rc;
}

int gasma_greedy_batch_packed(gasma_ctx* ctx, const uint16_t* read_p0, const uint16_t* read_p1, const uint32_t* read_off,
                              const uint16_t* ref_p0, const uint16_t* ref_p1, const uint32_t* ref_off, int64_t n,
                              int32_t* cost, gasma_cigar_op* ops, int64_t ops_cap, uint32_t* op_off) {
    if (!ctx) return GASMA_E_ARG;
    if (n < 0 || !read_off || !ref_off || !cost || !read_p0 || !read_p1 || !ref_p0 || !ref_p1) return fail(ctx, GASMA_E_ARG, "bad argument");
    if ((ops != nullptr) != (op_off != nullptr)) return fail(ctx, GASMA_E_ARG, "ops and op_off must be given together");
    if (ctx->W > 4) return fail(ctx, GASMA_E_UNSUPPORTED, "packed input covers max_length <= 256 (one alignment per thread)");
    if (n == 0) { if (op_off) op_off[0] = 0; return GASMA_OK; }
    int rc;
    const bool pipelined = n >= 65536;
    if (!pipelined) {
        if ((rc = check_offsets(ctx, read_off, n, "read"))) return rc;
        if ((rc = check_offsets(ctx, ref_off, n, "ref"))) return rc;
    }
    CU(cudaSetDevice(ctx->device));
    const size_t rb = 2 * (((size_t)read_off[n] + 15) >> 4), fb = 2 * (((size_t)ref_off[n] + 15) >> 4), ob = (size_t)(n + 1) * 4;
    if ((rc = reserve(ctx, ctx->d_reads, rb + 16))) return rc;
    if ((rc = reserve(ctx, ctx->d_r1, rb + 16))) return rc;
    if ((rc = reserve(ctx, ctx->d_refs, fb + 16))) return rc;
    if ((rc = reserve(ctx, ctx->d_f1, fb + 16))) return rc;
    if ((rc = reserve(ctx, ctx->d_read_off, ob))) return rc;
    if ((rc = reserve(ctx, ctx->d_ref_off, ob))) return rc;
    if ((rc = reserve(ctx, ctx->d_cost, (size_t)n * 4))) return rc;
    if (ops) {
        if ((rc = reserve(ctx, ctx->d_op_off, ob))) return rc;
        if ((rc = reserve(ctx, ctx->d_ops, (size_t)(ops_cap > 0 ? ops_cap : 1) * 2))) return rc;
    }
    if (pipelined)
        return greedy_batch_pipelined(ctx, (const char*)read_p0, read_off, (const char*)ref_p0, ref_off, n, cost, ops, ops_cap, op_off, read_p1, ref_p1);
    if (rb) { CU(cudaMemcpyAsync(ctx->d_reads.p, read_p0, rb, cudaMemcpyHostToDevice, ctx->stream)); CU(cudaMemcpyAsync(ctx->d_r1.p, read_p1, rb, cudaMemcpyHostToDevice, ctx->stream)); }
    if (fb) { CU(cudaMemcpyAsync(ctx->d_refs.p, ref_p0, fb, cudaMemcpyHostToDevice, ctx->stream)); CU(cudaMemcpyAsync(ctx->d_f1.p, ref_p1, fb, cudaMemcpyHostToDevice, ctx->stream)); }
    CU(cudaMemcpyAsync(ctx->d_read_off.p, read_off, ob, cudaMemcpyHostToDevice, ctx->stream));
    CU(cudaMemcpyAsync(ctx->d_ref_off.p, ref_off, ob, cudaMemcpyHostToDevice, ctx->stream));
    rc = gasma_greedy_batch_packed_dev(ctx, (const uint16_t*)ctx->d_reads.p, (const uint16_t*)ctx->d_r1.p, (const uint32_t*)ctx->d_read_off.p,
                                       (const uint16_t*)ctx->d_refs.p, (const uint16_t*)ctx->d_f1.p, (const uint32_t*)ctx->d_ref_off.p, n,
                                       (int32_t*)ctx->d_cost.p, ops ? (gasma_cigar_op*)ctx->d_ops.p : nullptr, ops_cap,
                                       ops ? (uint32_t*)ctx->d_op_off.p : nullptr);
    if (rc) return rc;
    CU(cudaMemcpyAsync(cost, ctx->d_cost.p, (size_t)n * 4, cudaMemcpyDeviceToHost, ctx->stream));
    if (ops) CU(cudaMemcpyAsync(op_off, ctx->d_op_off.p, ob, cudaMemcpyDeviceToHost, ctx->stream));
    if ((rc = gasma_synchronize(ctx))) return rc;
    if (ops) {
        const int64_t total = op_off[n];
        if (total > ops_cap) return fail(ctx, GASMA_E_CAPACITY, "ops buffer holds %lld ops, %lld needed", (long long)ops_cap, (long long)total);
        if (total) CU(cudaMemcpyAsync(ops, ctx->d_ops.p, (size_t)total * 2, cudaMemcpyDeviceToHost, ctx->stream));
        CU(cudaStreamSynchronize(ctx->stream));
    }
    return GASMA_OK;
}

}  // extern "C"

// =================================================================================================
// synthetic inputs
// =================================================================================================
extern "C" {

int64_t gasma_generate_ref_capacity(const gasma_gen_config* cfg, int64_t n_pairs) {
    if (!cfg || n_pairs < 0) return -1;
    GenCfg c{cfg->seed, cfg->length, cfg->mode, cfg->num_errors, cfg->thr_mismatch, cfg->thr_sub, cfg->thr_ins, cfg->thr_del};
    return (int64_t)gen_ref_capacity(c) * n_pairs + 16;
}

int gasma_generate_dev(gasma_ctx* ctx, const gasma_gen_config* cfg, uint64_t first_pair, int64_t n,
                       char* d_reads, uint32_t* d_read_off, char* d_refs, uint32_t* d_ref_off) {
    if (!ctx) return GASMA_E_ARG;
    if (!cfg || n < 0 || !d_reads || !d_read_off || !d_refs || !d_ref_off) return fail(ctx, GASMA_E_ARG, "bad argument");
    if (cfg->length < 0 || cfg->length > 1024 || cfg->num_errors < 0 || cfg->num_errors > 1024 || cfg->mode < 0 || cfg->mode > 2)
        return fail(ctx, GASMA_E_ARG, "generator: length/num_errors must be in [0,1024], mode in {0,1,2}");
    if ((int64_t)gasma_generate_ref_capacity(cfg, n) >= ((int64_t)1 << 32))
        return fail(ctx, GASMA_E_ARG, "generator: batch exceeds 4 GiB of bases");
    CU(cudaSetDevice(ctx->device));
    if (n == 0) { CU(cudaMemsetAsync(d_read_off, 0, 4, ctx->stream)); CU(cudaMemsetAsync(d_ref_off, 0, 4, ctx->stream)); return GASMA_OK; }
    GenCfg c{cfg->seed, cfg->length, cfg->mode, cfg->num_errors, cfg->thr_mismatch, cfg->thr_sub, cfg->thr_ins, cfg->thr_del};
    const int T = 128;
    const int64_t ntiles = (n + T - 1) / T;
    int rc;
    if ((rc = reserve(ctx, ctx->gen_len, (size_t)n * 4))) return rc;
    if ((rc = reserve(ctx, ctx->tile_sums, (size_t)ntiles * 4))) return rc;
    if ((rc = reserve(ctx, ctx->tile_bases, (size_t)ntiles * 4))) return rc;
    gen_lengths_kernel<T><<<(unsigned)ntiles, T, 0, ctx->stream>>>(c, first_pair, (long long)n, (uint32_t*)ctx->gen_len.p,
                                                                    (uint32_t*)ctx->tile_sums.p, d_read_off);
    CU(cudaGetLastError());
    scan_tiles_kernel<<<1, 1024, 0, ctx->stream>>>((uint32_t*)ctx->tile_sums.p, (uint32_t*)ctx->tile_bases.p, (int)ntiles, nullptr, nullptr);
    CU(cudaGetLastError());
    gen_write_kernel<T><<<(unsigned)ntiles, T, 0, ctx->stream>>>(c, first_pair, (long long)n, (uint32_t*)ctx->gen_len.p,
                                                                  (uint32_t*)ctx->tile_bases.p, d_reads, d_refs, d_ref_off);
    CU(cudaGetLastError());
    ctx->launches += 3;
    return GASMA_OK;
}

}  // extern "C"

// =================================================================================================
// K4 / K5 entry points (nw.cuh / metrics.cuh)
// =================================================================================================
#include "nw_entry.inl"
#include "leap_entry.inl"
#include "bench_entry.inl"
#include "noalloc_entry.inl"
#include "nccl_entry.inl"
#include "async_entry.inl"
#if GASMA_TIMELINE
extern "C" int gasma_debug_timeline(unsigned long long* out) { return (int)cudaMemcpyFromSymbol(out, gasma::g_tl, sizeof(unsigned long long) * 4 * 8192); }
#endif
