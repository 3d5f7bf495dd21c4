/* gasma_b200.h -- C ABI of libgasma_b200.so: the B200 (sm_100a) implementation of GASMA's
 * greedy bit-parallel alignment hot path.
 *
 * This is the drop-in boundary.  Every entry point is plain C: pointers,
 * sizes, one opaque context.  Each one names the reference interface it replaces; file:line
 * citations are relative to the reference tree (GZHoffie/approximate-string-matching).
 * The header-only C++ adapters in approximate-string-matching_b200/host/ put the reference's own
 * class names (hurdle_matrix, benchmark) back on top of these calls; INTEGRATION.md shows the
 * binding a maintainer of the reference would add.
 *
 * Conventions
 *   * Batches are "CSR strings": one concatenated ASCII buffer + (n+1) uint32 offsets; string i is
 *     bytes [off[i], off[i+1]).  A batch therefore holds < 4 GiB of bases per side.
 *   * Results are bit-exact to the reference's CPU output in canonical mode (staging tails = 'A',
 *     no FMA contraction).  Strings longer than max_length are rejected
 *     (GASMA_E_LENGTH) -- the C++ adapter applies the reference's silent truncation
 *     (hurdle_matrix.h:626-627) before calling.
 *   * Every function returns 0 (GASMA_OK) or a negative gasma_status; gasma_last_error() gives text.
 *   * "_dev" variants take DEVICE pointers (inputs resident in HBM, results left in HBM) and only
 *     enqueue work on the context's stream; the host variants take HOST pointers, copy in/out and
 *     synchronise before returning.
 *   * A context is single-caller (like one hurdle_matrix object) and is bound to one CUDA device.
 *     There is no CPU fallback: without a usable GPU gasma_create fails with GASMA_E_CUDA.
 */
#ifndef GASMA_B200_H
#define GASMA_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define GASMA_MAX_K 50          /* hurdle_matrix.h:8   MAX_K */
#define GASMA_WORD_BITS 64
#define GASMA_MAX_WORDS 16      /* max_length <= 1024 (reference: 128, utils.h:23-25); NW verifier: <= 256 */

typedef enum {
    GASMA_OK = 0,
    GASMA_E_ARG = -1,        /* null pointer, bad size, unaligned device buffer */
    GASMA_E_LENGTH = -2,     /* a string is longer than max_length */
    GASMA_E_BAND = -3,       /* k > GASMA_MAX_K, or |ref_len - read_len| > GASMA_MAX_K (out of bounds in the reference) */
    GASMA_E_CAPACITY = -4,   /* caller's ops buffer too small (op_off[n] tells the size needed) */
    GASMA_E_CUDA = -5,       /* CUDA runtime / launch failure, or no sm_100 device */
    GASMA_E_UNSUPPORTED = -6,/* parameter combination outside what the kernels cover */
    GASMA_E_INTERNAL = -7    /* step limit hit (non-terminating traversal) */
} gasma_status;

/* alignment_type_t, utils.h:554 */
typedef enum { GASMA_GLOBAL = 0, GASMA_SEMI_GLOBAL = 1 } gasma_alignment_type;

/* Constructor arguments of hurdle_matrix<T> (hurdle_matrix.h:552-562) + the per-reset band k (:625)
 * + the bit-vector width.  max_length = 64*W; 128 reproduces hurdle_matrix<int_128bit>. */
typedef struct {
    int32_t alignment_type;   /* gasma_alignment_type */
    int32_t x, o, e;          /* mismatch, gap open, gap extend (hurdle_matrix.h:478-480) */
    int32_t k;                /* band: lanes -k..k (hurdle_matrix.h:643-644) */
    int32_t max_length;       /* 64, 128, 192, 256: one alignment per thread; 512, 1024: one alignment per warp */
    double match_prob, mismatch_prob, indel_prob; /* hurdle_matrix.h:557-559 */
} gasma_params;

/* fills the reference's defaults: GLOBAL, x=o=e=1, k=3, max_length=128, probs 0.80, 0.20/3, 0.40/3 */
void gasma_default_params(gasma_params* p);

/* CIGAR op = (length << 3) | code.  Ops are never merged (hurdle_matrix.h:238-251 appends per step). */
typedef uint16_t gasma_cigar_op;
enum { GASMA_OP_M = 0, GASMA_OP_I = 1, GASMA_OP_D = 2, GASMA_OP_EQ = 3, GASMA_OP_X = 4 };
#define GASMA_OP_LEN(op) ((op) >> 3)
#define GASMA_OP_CODE(op) ((op) & 7)

typedef struct gasma_ctx gasma_ctx;

/* -- lifetime -------------------------------------------------------------------------------- */
/* Replaces `new hurdle_matrix<int_128bit>(GLOBAL,x,o,e)` (benchmark_utils.h:287). */
int gasma_create(const gasma_params* params, int device, gasma_ctx** out);
void gasma_destroy(gasma_ctx* ctx);
/* Text of the last failure on this context (or of the last failed gasma_create when ctx==NULL). */
const char* gasma_last_error(const gasma_ctx* ctx);
/* Change the band between batches: the `error` argument of reset() (hurdle_matrix.h:625). */
int gasma_set_band(gasma_ctx* ctx, int k);
/* The reference's compile-time band variant `#ifdef CORRECTION` (hurdle_matrix.h:501-512, :634-641) as a run-time
 * switch: with enabled != 0 the lanes of a pair are [-k, n-m+k] when m <= n and [n-m-k, k] otherwise (n-m =
 * ref_len - read_len), so the destination lane is always inside the band.  max_delta is the largest |n-m| the
 * context keeps lane slots for (2k+1+max_delta lanes per alignment in shared memory); a pair beyond it, or with
 * |n-m| + k > GASMA_MAX_K (past the reference's lane arrays), fails the batch with GASMA_E_BAND.  Results are
 * bit-exact to the unmodified reference compiled with -DCORRECTION.  enabled == 0 restores lanes -k..k. */
int gasma_set_correction(gasma_ctx* ctx, int enabled, int max_delta);
/* The aligner's other mask variants.  hurdle_matrix.h:453 runs the highway search on
 * mask.flip_short_hurdles(1) and carries a commented-out .flip_short_matches(1); int_128bit also offers
 * flip_short_hurdles(2) (utils.h:200-240).  variant 0 = flip_short_hurdles(1) (the reference as shipped, default),
 * 1 = flip_short_hurdles(2), 2 = flip_short_hurdles(1).flip_short_matches(1).  Penalties and hurdle counts are always
 * taken from the original mask, as in the reference.  Bit-exact to the unmodified reference with that call site
 * changed (oracle/ref_driver.cpp does it with a macro around the unmodified header).  max_length <= 256; not combined
 * with gasma_set_correction.  The int_256bit type (utils.h:274-549, README "Using AVX2") is NOT offered: in the
 * reference it reads an uninitialised upper half (hurdle_matrix.h:262-275), so it has no defined result. */
int gasma_set_mask_variant(gasma_ctx* ctx, int variant);
/* Use the caller's cudaStream_t (e.g. torch's current stream) for all work of this context.  The handle is
 * taken literally: NULL means the legacy default stream.  A fresh context uses a private non-blocking stream. */
int gasma_set_stream(gasma_ctx* ctx, void* cuda_stream);
/* Blocks until everything enqueued on the context's stream has finished; surfaces async errors
 * and the device-side status word (length / band / step-limit violations seen by _dev calls). */
int gasma_synchronize(gasma_ctx* ctx);
/* Number of kernels this context has launched so far (bench.py's gpu_launches). */
int64_t gasma_launch_count(const gasma_ctx* ctx);
/* Measurement hook for bench.py: summed device time (CUDA events recorded on the context's stream around
 * every launch) of the dominant kernel since the previous call -- which = 0 the greedy kernel (K2+K3),
 * 1 the NW kernel (K4), 2 the LEAP comparator kernel.  *launches receives how many launches the sum covers (the last 64 at most).
 * Blocks until those launches have finished. */
double gasma_kernel_time_ms(gasma_ctx* ctx, int which, int64_t* launches);
/* Sum of _step() calls (hurdle_matrix.h:407) over all pairs since the last call; resets the counter. */
int64_t gasma_last_step_count(gasma_ctx* ctx);

/* -- K1: 2-bit packing ------------------------------------------------------------------------- */
/* Replaces sse3_convert2bit1 (bit_convert.cpp:248, bit_convert.h:21) applied to a whole batch:
 * plane0[i*W + w] bit b = (char 64w+b of string i is C or T), plane1 likewise for G or T, anything
 * else (and the tail beyond the string) 0.  W = max_length/64. */
int gasma_pack2bit_batch(gasma_ctx* ctx, const char* seqs, const uint32_t* off, int64_t n,
                         uint64_t* plane0, uint64_t* plane1);
int gasma_pack2bit_batch_dev(gasma_ctx* ctx, const char* d_seqs, const uint32_t* d_off, int64_t n,
                             uint64_t* d_plane0, uint64_t* d_plane1);

/* -- K2+K3: greedy alignment ------------------------------------------------------------------- */
/* Replaces the loop body of benchmark::_run_greedy (benchmark_utils.h:185-201):
 * reset(read,m,ref,n,k); run(); get_cost(); get_CIGAR()  (hurdle_matrix.h:625,568,677,613) for n pairs.
 *   cost[i]            = get_cost()
 *   ops[op_off[i] .. op_off[i+1])  = get_CIGAR() as encoded ops (render with gasma_render_cigar)
 * ops may be NULL (costs only; op_off is then not written either).  ops_cap is the capacity of ops in
 * elements; if it is too small the call fails with GASMA_E_CAPACITY after writing cost[] and op_off[]. */
int gasma_greedy_batch(gasma_ctx* ctx,
                       const char* reads, const uint32_t* read_off,
                       const char* refs, const uint32_t* ref_off, int64_t n_pairs,
                       int32_t* cost, gasma_cigar_op* ops, int64_t ops_cap, uint32_t* op_off);
int gasma_greedy_batch_dev(gasma_ctx* ctx,
                           const char* d_reads, const uint32_t* d_read_off,
                           const char* d_refs, const uint32_t* d_ref_off, int64_t n_pairs,
                           int32_t* d_cost, gasma_cigar_op* d_ops, int64_t ops_cap, uint32_t* d_op_off);

/* -- K2+K3 on pre-packed input ("2-bit streams") -------------------------------------------------- */
/* The shape of the reference's own batch driver LEAP_SIMD/main.cpp:108-140, which converts its reads to 2-bit form
 * BEFORE its timed loop: a caller that keeps reads packed moves 4x fewer bytes to the GPU (about 83 B per 150 bp pair
 * instead of 308 B of ASCII) and the kernel skips K1.
 * Packed form of a CSR string buffer: two bit streams with ONE BIT PER CHARACTER at the character's CSR position --
 * plane0 bit i = (character i is C or T), plane1 bit i = (G or T), anything else 00 (sse3_convert2bit1's code,
 * bit_convert.cpp:248) -- stored as uint16 words, word c = characters [16c, 16c+16), bit b = character 16c+b
 * (LSB first).  ceil(total_chars / 16) words per plane; the CSR offsets stay the character offsets.
 * gasma_pack2bit_stream converts a buffer; gasma_greedy_batch_packed is gasma_greedy_batch on such input (same
 * outputs, bit for bit).  max_length <= 256. */
int gasma_pack2bit_stream(gasma_ctx* ctx, const char* seqs, int64_t total_chars, uint16_t* plane0, uint16_t* plane1);
int gasma_pack2bit_stream_dev(gasma_ctx* ctx, const char* d_seqs, int64_t total_chars, uint16_t* d_plane0, uint16_t* d_plane1);
int gasma_greedy_batch_packed(gasma_ctx* ctx,
                              const uint16_t* read_plane0, const uint16_t* read_plane1, const uint32_t* read_off,
                              const uint16_t* ref_plane0, const uint16_t* ref_plane1, const uint32_t* ref_off, int64_t n_pairs,
                              int32_t* cost, gasma_cigar_op* ops, int64_t ops_cap, uint32_t* op_off);
int gasma_greedy_batch_packed_dev(gasma_ctx* ctx,
                                  const uint16_t* d_read_plane0, const uint16_t* d_read_plane1, const uint32_t* d_read_off,
                                  const uint16_t* d_ref_plane0, const uint16_t* d_ref_plane1, const uint32_t* d_ref_off, int64_t n_pairs,
                                  int32_t* d_cost, gasma_cigar_op* d_ops, int64_t ops_cap, uint32_t* d_op_off);

/* -- the host-pointer greedy call as an asynchronous pair ---------------------------------------- */
/* The reference's loop is synchronous (benchmark_utils.h:238-252: reset, run, get_cost per pair); a caller that
 * streams batches through a GPU wants the link busy all the time, and a synchronous batch call ends with a tail
 * (last kernels, results back, one host round trip for the CIGAR fetch) during which nothing is copied in.
 * gasma_greedy_submit / gasma_greedy_submit_packed start gasma_greedy_batch / gasma_greedy_batch_packed on the
 * given arguments and return at once with a ticket; gasma_greedy_wait(ticket) returns that call's status when it
 * has ended (results are in the caller's buffers then, bit for bit those of the synchronous call).  Up to TWO calls
 * may be in flight per context (each has its own device mirrors and streams: the context's device memory for
 * mirrors doubles), a third submit fails with GASMA_E_UNSUPPORTED until one has been waited for.  All buffers of a
 * submitted call, inputs included, belong to the library until its wait returns; pinned host memory keeps the
 * copies asynchronous.  The context's band / widened-band / mask-variant settings at submit time apply.
 * gasma_destroy waits for calls still in flight.  Synchronous calls on the same context may be mixed in freely. */
int gasma_greedy_submit(gasma_ctx* ctx,
                        const char* reads, const uint32_t* read_off,
                        const char* refs, const uint32_t* ref_off, int64_t n_pairs,
                        int32_t* cost, gasma_cigar_op* ops, int64_t ops_cap, uint32_t* op_off, int* ticket);
int gasma_greedy_submit_packed(gasma_ctx* ctx,
                               const uint16_t* read_plane0, const uint16_t* read_plane1, const uint32_t* read_off,
                               const uint16_t* ref_plane0, const uint16_t* ref_plane1, const uint32_t* ref_off, int64_t n_pairs,
                               int32_t* cost, gasma_cigar_op* ops, int64_t ops_cap, uint32_t* op_off, int* ticket);
int gasma_greedy_wait(gasma_ctx* ctx, int ticket);

/* -- the greedy step without any allocation ------------------------------------------------------- */
/* No counterpart in the reference; added for callers that cannot call
 * cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  Same inputs, outputs
 * and results as gasma_greedy_batch_dev, but the call allocates NOTHING: its scratch (per-pair CIGAR staging rows,
 * op counts, per-tile sums, chunk counter) is a file-scope __device__ array of the library, 8.8 MB that every GPU
 * which loads the library holds until the process exits.  Limits: n_pairs <= GASMA_NOALLOC_MAX_PAIRS and the
 * context's max_length <= GASMA_NOALLOC_MAX_LENGTH; anything larger is rejected with GASMA_E_UNSUPPORTED.
 * All calls on one device share the array: calls that overlap on the GPU (different streams or contexts) race on
 * it, so serialise them.  The array keeps the previous call's contents; the call clears what must start at zero. */
#define GASMA_NOALLOC_MAX_PAIRS 8192
#define GASMA_NOALLOC_MAX_LENGTH 256
int gasma_greedy_batch_dev_noalloc(gasma_ctx* ctx,
                                   const char* d_reads, const uint32_t* d_read_off,
                                   const char* d_refs, const uint32_t* d_ref_off, int64_t n_pairs,
                                   int32_t* d_cost, gasma_cigar_op* d_ops, int64_t ops_cap, uint32_t* d_op_off);

/* -- LEAP comparator ----------------------------------------------------------- */
/* Replaces benchmark::_run_LEAP (GASMA/benchmark/benchmark_utils.h:156-178): class LV of
 * GASMA/benchmark/LEAP_SIMD/LV_BAG.cpp set up as init(k, 200, ED_GLOBAL, x, o, e) (:289); k, x, o, e are the
 * context's.  penalty[i] = LV::get_ED() after run() -- the only LEAP output the report uses (its accuracy line)
 * -- or INT32_MIN when no lane converges within 200 error levels (the reference then repeats the previous pair's
 * value).  Needs 1 <= e <= o, x and o <= 7, max_length <= 256; characters outside {A,C,G,T} count as 'A'. */
int gasma_leap_batch(gasma_ctx* ctx,
                     const char* reads, const uint32_t* read_off,
                     const char* refs, const uint32_t* ref_off, int64_t n_pairs, int32_t* penalty);
int gasma_leap_batch_dev(gasma_ctx* ctx,
                         const char* d_reads, const uint32_t* d_read_off,
                         const char* d_refs, const uint32_t* d_ref_off, int64_t n_pairs, int32_t* d_penalty);

/* -- K4: NW verifier ---------------------------------------------------------------------------- */
/* Replaces benchmark::_run_nw / _run_nw_sse (benchmark_utils.h:104-150; parasail_nw_trace* with
 * parasail_matrix_create("ACGT",0,-x), :288): optimal global affine-gap penalty (>= 0), gap of length g
 * costs o+(g-1)e.  d_bound may hold a per-pair upper bound on the optimum (e.g. the greedy cost) used
 * to size the band (exact); NULL = full matrix.  ops (optional) receive the traceback
 * as =,X,I,D runs. */
int gasma_nw_batch(gasma_ctx* ctx,
                   const char* reads, const uint32_t* read_off,
                   const char* refs, const uint32_t* ref_off, int64_t n_pairs,
                   const int32_t* bound, int32_t* penalty,
                   gasma_cigar_op* ops, int64_t ops_cap, uint32_t* op_off);
int gasma_nw_batch_dev(gasma_ctx* ctx,
                       const char* d_reads, const uint32_t* d_read_off,
                       const char* d_refs, const uint32_t* d_ref_off, int64_t n_pairs,
                       const int32_t* d_bound, int32_t* d_penalty,
                       gasma_cigar_op* d_ops, int64_t ops_cap, uint32_t* d_op_off);

/* -- K5: accuracy / coverage counters ----------------------------------------------------------- */
/* Replaces the counter block of benchmark::_run_benchmark (benchmark_utils.h:238,249-258) and
 * _check_coverage (:214-225; benchmark_coverage.h:26-91):
 *   counters[0] += n, [1] += #(nw == correct), [2] += #(greedy == correct), [3] += #(greedy LCM(>=1)
 *   covers NW LCM(>=3)).  answers may be NULL (correct = nw penalty, benchmark_utils.h:245-248).
 * Counters ACCUMULATE into d_counters (4 x uint64 in device memory) so that a multi-batch run ends
 * with one sum per GPU, which the caller all-reduces (NCCL) across GPUs. */
int gasma_metrics_dev(gasma_ctx* ctx, const char* d_reads, const uint32_t* d_read_off, int64_t n_pairs,
                      const int32_t* d_greedy_cost, const gasma_cigar_op* d_greedy_ops, const uint32_t* d_greedy_off,
                      const int32_t* d_nw_penalty, const gasma_cigar_op* d_nw_ops, const uint32_t* d_nw_off,
                      const int32_t* d_answers, uint64_t* d_counters);

/* -- the path's only collective -------------------------------------------------------------------- */
/* Sums the counters of a run that was sharded over several GPUs of one box (benchmark_utils.h:238,253-258: total,
 * nw_correct, greedy_correct, greedy_coverage): d_counters[i] = `count` uint64 words in the memory of ctxs[i]'s device
 * (what gasma_benchmark_batch_dev / gasma_metrics_dev accumulate into); after the call EVERY array holds the sum over
 * all contexts.  One ncclAllReduce(ncclSum, ncclUint64) per rank over NVLink / NVSwitch, enqueued on each context's
 * stream behind the kernels that produced its counts; the call returns when all ranks are done.  One context per
 * device; the communicators are built once per device list (ncclCommInitAll) and kept for the process.  NCCL is
 * loaded at run time (libnccl.so.2, or $GASMA_NCCL_LIB): GASMA_E_UNSUPPORTED if it is not installed -- the sum of
 * n_ctx x count integers on the host is then an exact substitute.  gasma_nccl_version() = NCCL's version code, 0 if
 * it cannot be loaded. */
int gasma_allreduce_counters(gasma_ctx* const* ctxs, int n_ctx, uint64_t* const* d_counters, int count);
/* The same for counts held in HOST arrays (C++ callers without CUDA headers, host/benchmark_gpu.h): each array goes to
 * its context's device, the device arrays are all-reduced, every host array comes back holding the total. */
int gasma_allreduce_counters_host(gasma_ctx* const* ctxs, int n_ctx, uint64_t* const* counters, int count);
int gasma_nccl_version(void);

/* -- the whole benchmark body, batched ------------------------------------------------------------ */
/* Replaces benchmark::run's loop over _run_benchmark (benchmark_utils.h:373-385, :231-259) for n pairs:
 * greedy (reset/run/get_cost/get_CIGAR), NW verifier (banded by the greedy penalty, exact), and the four
 * counters, entirely on the device; only the counters (and the optional per-pair penalties) come back.
 *   counters[0..3] += total, nw_correct, greedy_correct, greedy_coverage   (ACCUMULATES, like the class members)
 *   answers     optional n x int32 (read_answer_file, :358-368); INT32_MIN or NULL = use the NW penalty
 *   greedy_cost, nw_penalty   optional n x int32 outputs
 *   device_ms   optional: [0] += greedy kernels, [1] += NW kernels (CUDA-event time, the report's [Time] rows)
 * The LEAP column of the report (LV_BAG) is a comparator, not part of this path, and is not computed. */
int gasma_benchmark_batch(gasma_ctx* ctx,
                          const char* reads, const uint32_t* read_off,
                          const char* refs, const uint32_t* ref_off, int64_t n_pairs,
                          const int32_t* answers, uint64_t counters[4],
                          int32_t* greedy_cost, int32_t* nw_penalty, double device_ms[2]);
/* Same with the batch resident in HBM; d_counters = 4 x uint64 in device memory (accumulated; all-reduce it
 * across GPUs at the end of a run).  d_greedy_cost / d_nw_penalty are optional. */
int gasma_benchmark_batch_dev(gasma_ctx* ctx,
                              const char* d_reads, const uint32_t* d_read_off,
                              const char* d_refs, const uint32_t* d_ref_off, int64_t n_pairs,
                              const int32_t* d_answers, uint64_t* d_counters,
                              int32_t* d_greedy_cost, int32_t* d_nw_penalty);

/* -- synthetic inputs ---------------------------------------------------------------------------- */
/* Restates Dataset (benchmark_dataset.h:100-186) with a counter RNG keyed by (seed, pair index) so
 * every shard regenerates pair i identically.  mode 0: exactly num_errors edits; 1: uniform in
 * [0,num_errors); 2: per-base profile (thr_* are cumulative thresholds on a 32-bit draw). */
typedef struct {
    uint64_t seed;
    int32_t length, mode, num_errors, reserved;
    uint64_t thr_mismatch, thr_sub, thr_ins, thr_del;
} gasma_gen_config;
/* bytes of ref buffer that n pairs can need at most */
int64_t gasma_generate_ref_capacity(const gasma_gen_config* cfg, int64_t n_pairs);
/* Writes reads (n*length bytes), read_off, refs, ref_off (n+1 each) for pairs first..first+n. */
int gasma_generate_dev(gasma_ctx* ctx, const gasma_gen_config* cfg, uint64_t first_pair, int64_t n_pairs,
                       char* d_reads, uint32_t* d_read_off, char* d_refs, uint32_t* d_ref_off);

/* Host-pointer variant: generates on the device and copies the batch out (reads: n*length bytes, refs:
 * gasma_generate_ref_capacity bytes, offsets n+1 each). */
int gasma_generate(gasma_ctx* ctx, const gasma_gen_config* cfg, uint64_t first_pair, int64_t n_pairs,
                   char* reads, uint32_t* read_off, char* refs, uint32_t* ref_off);

/* -- host helpers -------------------------------------------------------------------------------- */
/* get_CIGAR() text from encoded ops: std::to_string(len) + letter per op (hurdle_matrix.h:241-249).
 * Returns the text length (without NUL); writes at most cap bytes including the NUL. */
int gasma_render_cigar(const gasma_cigar_op* ops, int n_ops, char* out, int cap);

#ifdef __cplusplus
}
#endif
#endif /* GASMA_B200_H */
