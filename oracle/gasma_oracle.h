/* oracle/gasma_oracle.h -- TEST INFRASTRUCTURE (the checker), NOT PRODUCT CODE.
 *
 * Plain-C CPU restatement of the reference's greedy-alignment hot path, generalised
 * from the reference's 128-bit vectors to W 64-bit words (W=2 is the reference's
 * own domain, hurdle_matrix<int_128bit>).  Only tests/, __graft_entry__.smoke() and
 * bench.py's cpu_baseline / --impl reference legs may load it.  The product
 * (libgasma_b200.so) never links or calls anything in oracle/.
 *
 * Parity pinning: at W=2 this file is checked against oracle/_ref/libgasma_ref.so
 * (the UNMODIFIED reference compiled in place, canonical mode) on the golden
 * vectors of SURVEY.md section 8c and on seeded bulk batches -- see
 * tests/test_oracle.py and tests/golden/.  The NW traceback (hence Coverage %)
 * is "parity unpinned": parasail is an un-vendored submodule of the reference
 * (jeffdaily/parasail @ b999b5eb), see DESIGN.md.
 *
 * All file:line citations are relative to /root/reference.
 */
#ifndef GASMA_ORACLE_H
#define GASMA_ORACLE_H
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define GO_MAX_W 16   /* 16 x 64 = 1024 bases */
#define GO_MAX_K 50   /* hurdle_matrix.h:8 */

/* CIGAR op = (length << 3) | code ; same encoding as include/gasma_b200.h */
enum { GO_OP_M = 0, GO_OP_I = 1, GO_OP_D = 2, GO_OP_EQ = 3, GO_OP_X = 4 };

typedef struct {
    int k, x, o, e;
    int global;          /* 1 = GLOBAL, 0 = SEMI_GLOBAL (utils.h:554) */
    int W;               /* words per bit-vector; MAX_LENGTH = 64*W */
    double match_prob, mismatch_prob, indel_prob; /* hurdle_matrix.h:557-559 */
    int correction;      /* 1 = the reference built with -DCORRECTION (hurdle_matrix.h:501-512, :634-641):
                          * lanes [-k, n-m+k] for m <= n, [n-m-k, k] otherwise; 0 = lanes -k..k */
    int mask_variant;    /* lane the highway search runs on (hurdle_matrix.h:453): 0 = flip_short_hurdles(1) as shipped,
                          * 1 = flip_short_hurdles(2), 2 = flip_short_hurdles(1).flip_short_matches(1) (utils.h:200-240) */
} go_params;

/* bit_convert.cpp:248 semantics for one string, tail [len, 64W) = 'A' (00) */
void go_pack2bit(const char* s, int len, int W, uint64_t* plane0, uint64_t* plane1);

/* One alignment: returns 0 ok.  ops (capacity cap) receives encoded ops, *nops the count
 * (count keeps growing past cap so the caller can detect overflow). */
int go_greedy(const char* read, int read_len, const char* ref, int ref_len, const go_params* p,
              int32_t* cost, uint16_t* ops, int cap, int* nops, int* nsteps);

/* Batch over CSR-laid-out strings. cigar_text may be NULL; otherwise stride bytes/pair, NUL terminated.
 * ops/op_off may be NULL; op_off has n+1 entries (CSR). threads>1 splits contiguous shards. */
int go_greedy_batch(const char* reads, const uint32_t* read_off, const char* refs, const uint32_t* ref_off,
                    int64_t n, const go_params* p, int threads,
                    int32_t* cost, char* cigar_text, int stride,
                    uint16_t* ops, uint32_t* op_off, int64_t ops_cap, int64_t* total_steps);

/* render ops -> text with std::to_string semantics (hurdle_matrix.h:238-251); returns length */
int go_render_cigar(const uint16_t* ops, int nops, char* out, int cap);

/* Scalar Gotoh global affine DP, parasail scoring convention (benchmark_utils.h:113-116,288):
 * match 0, mismatch x, gap of length g costs o+(g-1)e.  Returns the penalty (>=0).
 * ops (may be NULL) receives the traceback in forward order with codes = / X / I / D. */
int go_nw(const char* read, int m, const char* ref, int n, int x, int o, int e,
          uint16_t* ops, int cap, int* nops);
int go_nw_batch(const char* reads, const uint32_t* read_off, const char* refs, const uint32_t* ref_off,
                int64_t n, int x, int o, int e, int threads, int32_t* penalty,
                uint16_t* ops, uint32_t* op_off, int64_t ops_cap);

/* LEAP comparator (GASMA/benchmark/LEAP_SIMD/LV_BAG.cpp as driven by benchmark_utils.h:156-178,289):
 * the penalty the benchmark records for it, INT32_MIN if no lane converges within 200 errors. */
int go_leap(const char* read, int m, const char* ref, int n, int k, int x, int o, int e);
int go_leap_batch(const char* reads, const uint32_t* read_off, const char* refs, const uint32_t* ref_off,
                  int64_t n, int k, int x, int o, int e, int threads, int32_t* penalty);

/* benchmark_coverage.h:26-67 / :73-91 on encoded ops */
int go_lcm(const char* s1, int s1_len, const uint16_t* ops, int nops, int threshold, char* out, int cap);
int go_covers(const char* a, int na, const char* b, int nb);

/* benchmark_utils.h:231-259 counters: out[0]=total, [1]=nw_correct, [2]=greedy_correct, [3]=greedy_coverage */
void go_metrics(const char* reads, const uint32_t* read_off, int64_t n,
                const int32_t* greedy_cost, const uint16_t* g_ops, const uint32_t* g_off,
                const int32_t* nw_penalty, const uint16_t* n_ops, const uint32_t* n_off,
                const int32_t* answers /* may be NULL; INT32_MIN = use nw */, uint64_t out[4]);

/* Synthetic pair generator: CPU twin of the device generator (restates Dataset,
 * benchmark_dataset.h:100-186, with a counter RNG keyed by (seed, pair index)). */
typedef struct {
    uint64_t seed;
    int length;          /* read length L */
    int mode;            /* 0 = exactly num_errors edits, 1 = uniform in [0,num_errors), 2 = per-base profile */
    int num_errors;      /* ceil(L * e) computed by the caller in float like benchmark_dataset.h:153-156 */
    uint64_t thr_mismatch; /* edit is a mismatch iff (u32 draw) < thr ; = mismatch_rate * 2^32 */
    uint64_t thr_sub, thr_ins, thr_del; /* mode 2: per-base thresholds on a u32 draw (cumulative: del, ins, sub) */
} go_gen_config;
int go_gen_ref_capacity(const go_gen_config* c);
/* writes read (length L) and ref; returns ref length */
int go_generate_pair(const go_gen_config* c, uint64_t pair_index, char* read, char* ref);

#ifdef __cplusplus
}
#endif
#endif
