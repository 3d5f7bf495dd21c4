/* oracle/gasma_oracle.c -- TEST INFRASTRUCTURE (the checker), NOT PRODUCT CODE.
 * See gasma_oracle.h for the role of this file and how its parity is pinned.
 * Build with -ffp-contract=off (part of the parity contract, oracle/Makefile).
 * All file:line citations are relative to /root/reference.
 */
#define _GNU_SOURCE
#include "gasma_oracle.h"
#include <limits.h>
#include <math.h>
#include <pthread.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

/* ------------------------------------------------------------------------------------------
 * W-word little-endian bit vector; bit i <-> character i.  Semantics of int_128bit
 * (GASMA/utils.h:49-271) generalised to BITS = 64*W.  NB the reference's names are inverted
 * w.r.t. the bit index: its shift_left() moves towards bit 0 (utils.h:143-153).
 * ---------------------------------------------------------------------------------------- */
typedef struct { uint64_t w[GO_MAX_W]; } bv_t;

static void bv_zero(bv_t* v) { memset(v->w, 0, sizeof(v->w)); }

/* towards bit 0, zero fill, s >= BITS (or s < 0, see SSE count semantics) -> 0   utils.h:143 */
static bv_t bv_down(const bv_t* v, int s, int W) {
    bv_t r; bv_zero(&r);
    if (s < 0 || s >= 64 * W) return r;
    int ws = s >> 6, bs = s & 63;
    for (int i = 0; i + ws < W; ++i) {
        uint64_t lo = v->w[i + ws] >> bs;
        uint64_t hi = (bs && i + ws + 1 < W) ? v->w[i + ws + 1] << (64 - bs) : 0;
        r.w[i] = lo | hi;
    }
    return r;
}
/* away from bit 0, truncated to BITS   utils.h:131 */
static bv_t bv_up(const bv_t* v, int s, int W) {
    bv_t r; bv_zero(&r);
    if (s < 0 || s >= 64 * W) return r;
    int ws = s >> 6, bs = s & 63;
    for (int i = W - 1; i - ws >= 0; --i) {
        uint64_t hi = v->w[i - ws] << bs;
        uint64_t lo = (bs && i - ws - 1 >= 0) ? v->w[i - ws - 1] >> (64 - bs) : 0;
        r.w[i] = hi | lo;
    }
    return r;
}
/* utils.h:168-182: index of lowest set bit, BITS when empty */
static int bv_first_one(const bv_t* v, int W) {
    for (int i = 0; i < W; ++i)
        if (v->w[i]) return 64 * i + __builtin_ctzll(v->w[i]);
    return 64 * W;
}
/* utils.h:187-191 */
static int bv_first_zero(const bv_t* v, int W) {
    for (int i = 0; i < W; ++i)
        if (~v->w[i]) return 64 * i + __builtin_ctzll(~v->w[i]);
    return 64 * W;
}
static int bv_popcount(const bv_t* v, int W) {
    int c = 0;
    for (int i = 0; i < W; ++i) c += __builtin_popcountll(v->w[i]);
    return c;
}
/* utils.h:263-270: shift_left(from).shift_right(from + BITS - to).pop_count() */
static int bv_pcb(const bv_t* v, int from, int to, int W) {
    bv_t a = bv_down(v, from, W);
    bv_t b = bv_up(&a, from + 64 * W - to, W);
    return bv_popcount(&b, W);
}
/* utils.h:200-216: threshold 1: v & ((v>>1) | (v<<1)); threshold 2: v & ((v>>1) | (v<<1) | (v>>2) | (v<<2)) */
static bv_t bv_flip_short_hurdles(const bv_t* v, int W, int threshold) {
    bv_t l = bv_down(v, 1, W), r = bv_up(v, 1, W), out; bv_zero(&out);
    bv_t l2, r2; bv_zero(&l2); bv_zero(&r2);
    if (threshold > 1) { l2 = bv_down(v, 2, W); r2 = bv_up(v, 2, W); }
    for (int i = 0; i < W; ++i) out.w[i] = v->w[i] & (l.w[i] | r.w[i] | l2.w[i] | r2.w[i]);
    return out;
}
/* utils.h:223-240 with threshold 1: v | (shift_left_one(v) & shift_right_one(v)): a 0 between two 1s becomes 1;
 * shift_left_one (:160-163) feeds a 1 into the top bit, shift_right_one (:155-158) a 1 into bit 0 */
static bv_t bv_flip_short_matches(const bv_t* v, int W) {
    bv_t l = bv_down(v, 1, W), r = bv_up(v, 1, W), out; bv_zero(&out);
    l.w[W - 1] |= 1ull << 63;
    r.w[0] |= 1ull;
    for (int i = 0; i < W; ++i) out.w[i] = v->w[i] | (l.w[i] & r.w[i]);
    return out;
}
/* the lane the highway search runs on (hurdle_matrix.h:453 and its commented-out tail):
 * variant 0: flip_short_hurdles(1) (the reference as shipped); 1: flip_short_hurdles(2);
 * 2: flip_short_hurdles(1).flip_short_matches(1) */
static bv_t bv_search_lane(const bv_t* v, int W, int variant) {
    bv_t f = bv_flip_short_hurdles(v, W, variant == 1 ? 2 : 1);
    if (variant == 2) f = bv_flip_short_matches(&f, W);
    return f;
}

/* bit_convert.cpp:248-369 (output side only): plane0 bit i = c in {C,T}, plane1 bit i = c in {G,T},
 * everything else 00.  Tail [len, BITS) = 00 = 'A' (canonical mode). */
void go_pack2bit(const char* s, int len, int W, uint64_t* plane0, uint64_t* plane1) {
    for (int i = 0; i < W; ++i) plane0[i] = plane1[i] = 0;
    for (int i = 0; i < len && i < 64 * W; ++i) {
        char c = s[i];
        if (c == 'C' || c == 'T') plane0[i >> 6] |= 1ull << (i & 63);
        if (c == 'G' || c == 'T') plane1[i >> 6] |= 1ull << (i & 63);
    }
}

/* utils.h:576-579 */
static int slp(int l1, int l2, int o, int e) { return l1 == l2 ? 0 : o + e * (abs(l1 - l2) - 1); }
/* utils.h:587-593 */
static int sfc(int l1, int l2) {
    if (l1 * l2 >= 0) return abs(l1) > abs(l2) ? abs(l1) - abs(l2) : 0;
    return abs(l1);
}
/* hurdle_matrix.h:58-68 */
static int calc_destination(int m, int n, int lane) {
    if (m >= n) {
        if (lane > 0) return n - lane;
        else if (lane >= n - m) return n;
        else return m + lane;
    } else {
        if (lane < 0) return m + lane;
        else if (lane <= n - m) return m;
        else return n - lane;
    }
}

static void emit_op(uint16_t* ops, int cap, int* nops, int code, int len) {
    if (*nops < cap) ops[*nops] = (uint16_t)((len << 3) | code);
    ++*nops;
}
/* hurdle_matrix.h:238-251 */
static void update_cigar(uint16_t* ops, int cap, int* nops, int best, int cur, int total) {
    if (best < cur) emit_op(ops, cap, nops, GO_OP_I, cur - best);
    else if (best > cur) emit_op(ops, cap, nops, GO_OP_D, best - cur);
    if (total > 0) emit_op(ops, cap, nops, GO_OP_M, total);
}

int go_greedy(const char* read, int read_len, const char* ref, int ref_len, const go_params* p,
              int32_t* cost_out, uint16_t* ops, int cap, int* nops_out, int* nsteps_out) {
    const int W = p->W, BITS = 64 * W, k = p->k, x = p->x, o = p->o, e = p->e, GLOBAL = p->global;
    if (W < 1 || W > GO_MAX_W || k < 0 || k > GO_MAX_K) return -1;
    /* hurdle_matrix.h:626-627 silent truncation */
    const int m = read_len < BITS ? read_len : BITS, n = ref_len < BITS ? ref_len : BITS;
    if (abs(n - m) > GO_MAX_K) return -2;  /* out-of-bounds in the reference */
    /* lane range, hurdle_matrix.h:634-644 (#ifdef CORRECTION widens the band towards lane n-m) */
    int lb = -k, ub = k;
    if (p->correction) { if (m <= n) ub = n - m + k; else lb = n - m - k; }
    if (lb < -GO_MAX_K || ub > GO_MAX_K) return -2;  /* past the reference's lane arrays */
    /* hurdle_matrix.h:536-538 */
    const double ms = log(p->match_prob / 0.25), mms = log(p->mismatch_prob / 0.25),
                 ids = log(p->indel_prob / 2 / 0.25);

    bv_t A0, A1, B0, B1;
    bv_zero(&A0); bv_zero(&A1); bv_zero(&B0); bv_zero(&B1);
    go_pack2bit(read, m, W, A0.w, A1.w);
    go_pack2bit(ref, n, W, B0.w, B1.w);

    /* state of a fresh object: lanes default to value 1 (utils.h:57-59), destination computed with
     * m=n=0 (hurdle_matrix.h:79-92, ctor called with "" ""), starting_point -1 */
    static __thread bv_t orig[2 * GO_MAX_K + 1], flip[2 * GO_MAX_K + 1];
    int sp[2 * GO_MAX_K + 1], len[2 * GO_MAX_K + 1], sw[2 * GO_MAX_K + 1], hc[2 * GO_MAX_K + 1],
        nsw[2 * GO_MAX_K + 1], nh[2 * GO_MAX_K + 1], de[2 * GO_MAX_K + 1];
#define IX(l) ((l) + GO_MAX_K)
    for (int l = -GO_MAX_K; l <= GO_MAX_K; ++l) {
        if (l < lb || l > ub) {
            bv_zero(&orig[IX(l)]); orig[IX(l)].w[0] = 1;
            flip[IX(l)] = orig[IX(l)];
        }
        de[IX(l)] = calc_destination(0, 0, l);
        sp[IX(l)] = -1; len[IX(l)] = 0;
        sw[IX(l)] = hc[IX(l)] = nsw[IX(l)] = nh[IX(l)] = BITS;
    }
    /* _construct_hurdles, hurdle_matrix.h:441-455 ; highways::reset :106-119 */
    for (int l = lb; l <= ub; ++l) {
        bv_t s0, s1, v; bv_zero(&v);
        if (l < 0) {
            s0 = bv_down(&A0, -l, W); s1 = bv_down(&A1, -l, W);
            for (int i = 0; i < W; ++i) v.w[i] = (s0.w[i] ^ B0.w[i]) | (s1.w[i] ^ B1.w[i]);
        } else {
            s0 = bv_down(&B0, l, W); s1 = bv_down(&B1, l, W);
            for (int i = 0; i < W; ++i) v.w[i] = (s0.w[i] ^ A0.w[i]) | (s1.w[i] ^ A1.w[i]);
        }
        orig[IX(l)] = v;
        flip[IX(l)] = bv_search_lane(&v, W, p->mask_variant);
        sp[IX(l)] = -1;
        sw[IX(l)] = hc[IX(l)] = nsw[IX(l)] = nh[IX(l)] = BITS;
        de[IX(l)] = calc_destination(m, n, l);
    }
    const int dl = n - m;
    int cl = 0, cc = 0, cost = 0, first = 1, nops = 0, nsteps = 0;

    for (;;) {
        /* _update_highway_list loop 1, hurdle_matrix.h:291-322 */
        int reach = 0;
        for (int l = lb; l <= ub; ++l) {
            int sc = cc + sfc(cl, l);
            if (sp[IX(l)] < sc) {
                nsw[IX(l)] = abs(l - cl);
                bv_t t = bv_down(&flip[IX(l)], sc, W);
                int fz = bv_first_zero(&t, W);
                bv_t t2 = bv_down(&t, fz, W);
                int nx = bv_first_one(&t2, W);
                sp[IX(l)] = sc + fz;
                len[IX(l)] = nx;
                if (sc + fz + nx > de[IX(l)]) {
                    int d = de[IX(l)] - (sc + fz);
                    len[IX(l)] = d > 0 ? d : 0;
                    reach = 1;
                }
            }
            sw[IX(l)] = (GLOBAL || !first) ? slp(cl, l, o, e) : 0;
            nh[IX(l)] = bv_pcb(&orig[IX(l)], sc, sp[IX(l)] + len[IX(l)], W);
            hc[IX(l)] = x * nh[IX(l)];
        }
        /* loop 2, hurdle_matrix.h:325-352 */
        double bestH = -INFINITY;
        int bestLeap = 0; /* -numeric_limits<int>::infinity() == 0, hurdle_matrix.h:287 */
        int bl = 0;
        for (int l = lb; l <= ub; ++l) {
            int current_cost = -sw[IX(l)] - hc[IX(l)];
            double h = ms * len[IX(l)] + mms * nh[IX(l)] + ids * nsw[IX(l)];
            int lh = -sw[IX(l)];
            if (reach) {
                int f = GLOBAL ? slp(l, dl, o, e) : 0;
                h = current_cost - f - x * (de[IX(l)] - sp[IX(l)] - len[IX(l)]);
                lh -= f;
            }
            if (h > bestH || (h == bestH && lh > bestLeap)) { bestH = h; bestLeap = lh; bl = l; }
        }
        ++nsteps;
        if (len[IX(bl)] <= 0) break; /* :358-360 -> _step returns true */
        /* _choose_best_highway, hurdle_matrix.h:368-401 */
        int spb = sp[IX(bl)], sT = hc[IX(bl)] + sw[IX(bl)], sI = sT, b = bl;
        for (int l = lb; l <= ub; ++l) {
            if (l == bl) continue;
            if (sp[IX(l)] + sfc(l, bl) > spb) continue;
            int end = sp[IX(l)] + len[IX(l)];
            int ic = sw[IX(l)] + bv_pcb(&orig[IX(l)], cc + sfc(cl, l), end, W);
            int hp = x * bv_pcb(&orig[IX(bl)], sfc(l, bl) + end, spb, W);
            int tc = ic + slp(l, bl, o, e) + (hp > 0 ? hp : 0);
            if (tc <= sT && ic <= sI) { sT = tc; sI = ic; b = l; }
        }
        /* _step, hurdle_matrix.h:412-433 */
        cost += sw[IX(b)] + hc[IX(b)];
        int dist = sp[IX(b)] + len[IX(b)] - (cc + sfc(cl, b));
        update_cigar(ops, cap, &nops, b, cl, dist);
        cl = b;
        cc = sp[IX(b)] + len[IX(b)];
        first = 0;
        if (cc >= de[IX(cl)]) break;
    }
    /* run() fix-up, hurdle_matrix.h:575-590 */
    int dc = de[IX(dl)];
    if (cl != dl || cc < dc) {
        int d = bv_pcb(&orig[IX(dl)], cc + sfc(cl, dl), dc, W);
        int hcst = x * d;
        cost += (GLOBAL ? slp(cl, dl, o, e) : 0) + (hcst > 0 ? hcst : 0);
        update_cigar(ops, cap, &nops, dl, cl, d);
    }
#undef IX
    *cost_out = cost;
    if (nops_out) *nops_out = nops;
    if (nsteps_out) *nsteps_out = nsteps;
    return 0;
}

/* std::to_string + op letter, hurdle_matrix.h:241-249 */
int go_render_cigar(const uint16_t* ops, int nops, char* out, int cap) {
    static const char letters[] = "MID=X";
    int pos = 0;
    for (int i = 0; i < nops; ++i) {
        char tmp[16];
        int c = snprintf(tmp, sizeof tmp, "%d%c", ops[i] >> 3, letters[ops[i] & 7]);
        if (pos + c < cap) memcpy(out + pos, tmp, (size_t)c);
        pos += c;
    }
    if (cap > 0) out[pos < cap ? pos : cap - 1] = '\0';
    return pos;
}

/* ------------------------------------------------------------------------------------------
 * batch drivers
 * ---------------------------------------------------------------------------------------- */
typedef struct {
    const char* reads; const uint32_t* read_off; const char* refs; const uint32_t* ref_off;
    int64_t lo, hi; const go_params* p;
    int32_t* cost; char* cigar_text; int stride;
    uint16_t* ops; uint32_t* op_cnt; int64_t steps; int rc;
    int x, o, e; /* nw */
} shard_t;

#define GO_OPS_LOCAL 4096

static void* greedy_shard(void* arg) {
    shard_t* s = (shard_t*)arg;
    uint16_t local[GO_OPS_LOCAL];
    s->steps = 0; s->rc = 0;
    for (int64_t i = s->lo; i < s->hi; ++i) {
        int nops = 0, nsteps = 0; int32_t c = 0;
        int rc = go_greedy(s->reads + s->read_off[i], (int)(s->read_off[i + 1] - s->read_off[i]),
                           s->refs + s->ref_off[i], (int)(s->ref_off[i + 1] - s->ref_off[i]),
                           s->p, &c, local, GO_OPS_LOCAL, &nops, &nsteps);
        if (rc) { s->rc = rc; c = INT32_MIN; nops = 0; }
        s->steps += nsteps;
        if (s->cost) s->cost[i] = c;
        if (s->cigar_text) go_render_cigar(local, nops, s->cigar_text + (size_t)i * s->stride, s->stride);
        if (s->op_cnt) s->op_cnt[i] = (uint32_t)nops;
        /* ops are gathered in a second pass by the caller (needs the prefix sum) */
    }
    return NULL;
}

static void run_shards(void* (*fn)(void*), shard_t* proto, int64_t n, int threads, shard_t* out) {
    if (threads < 1) threads = 1;
    if (threads > 256) threads = 256;
    pthread_t th[256];
    for (int t = 0; t < threads; ++t) {
        out[t] = *proto;
        out[t].lo = n * t / threads; out[t].hi = n * (t + 1) / threads;
        if (threads == 1) fn(&out[t]); else pthread_create(&th[t], NULL, fn, &out[t]);
    }
    if (threads > 1) for (int t = 0; t < threads; ++t) pthread_join(th[t], NULL);
}

int go_greedy_batch(const char* reads, const uint32_t* read_off, const char* refs, const uint32_t* ref_off,
                    int64_t n, const go_params* p, int threads,
                    int32_t* cost, char* cigar_text, int stride,
                    uint16_t* ops, uint32_t* op_off, int64_t ops_cap, int64_t* total_steps) {
    shard_t proto; memset(&proto, 0, sizeof proto);
    proto.reads = reads; proto.read_off = read_off; proto.refs = refs; proto.ref_off = ref_off;
    proto.p = p; proto.cost = cost; proto.cigar_text = cigar_text; proto.stride = stride;
    uint32_t* cnt = op_off ? (uint32_t*)malloc(sizeof(uint32_t) * (size_t)(n + 1)) : NULL;
    proto.op_cnt = cnt;
    shard_t sh[256];
    if (threads > 256) threads = 256;
    run_shards(greedy_shard, &proto, n, threads, sh);
    int rc = 0; int64_t steps = 0;
    for (int t = 0; t < (threads < 1 ? 1 : threads); ++t) { if (sh[t].rc) rc = sh[t].rc; steps += sh[t].steps; }
    if (total_steps) *total_steps = steps;
    if (op_off) {
        uint32_t acc = 0;
        for (int64_t i = 0; i < n; ++i) { op_off[i] = acc; acc += cnt[i]; }
        op_off[n] = acc;
        free(cnt);
        if (ops) {
            if ((int64_t)acc > ops_cap) return -3;
            uint16_t local[GO_OPS_LOCAL];
            for (int64_t i = 0; i < n; ++i) { /* second pass: deterministic, single thread */
                int nops = 0; int32_t c;
                if (go_greedy(reads + read_off[i], (int)(read_off[i + 1] - read_off[i]), refs + ref_off[i],
                              (int)(ref_off[i + 1] - ref_off[i]), p, &c, local, GO_OPS_LOCAL, &nops, NULL) == 0)
                    memcpy(ops + op_off[i], local, sizeof(uint16_t) * (size_t)nops);
            }
        }
    }
    return rc;
}

/* ------------------------------------------------------------------------------------------
 * NW verifier: scalar Gotoh, global, affine.  Stands in for parasail_nw_trace*
 * (benchmark_utils.h:113-116,139-142; matrix "ACGT",0,-x at :288).  parasail convention: a gap
 * of length g costs o + (g-1)*e.  The optimum is unique, so the penalty is pinned mathematically;
 * the traceback rule below (diagonal, then I, then D; prefer closing a gap) is OUR choice --
 * "parity unpinned" for the CIGAR/coverage (see header).
 * ---------------------------------------------------------------------------------------- */
#define NW_INF (INT_MAX / 4)
int go_nw(const char* read, int m, const char* ref, int n, int x, int o, int e,
          uint16_t* ops, int cap, int* nops_out) {
    size_t cols = (size_t)n + 1, cells = ((size_t)m + 1) * cols;
    int* H = (int*)malloc(sizeof(int) * cells * 3);
    int* I = H + cells; int* D = I + cells;
#define AT(M, i, j) M[(size_t)(i) * cols + (size_t)(j)]
    AT(H, 0, 0) = 0; AT(I, 0, 0) = NW_INF; AT(D, 0, 0) = NW_INF;
    for (int j = 1; j <= n; ++j) { AT(D, 0, j) = o + (j - 1) * e; AT(H, 0, j) = AT(D, 0, j); AT(I, 0, j) = NW_INF; }
    for (int i = 1; i <= m; ++i) {
        AT(I, i, 0) = o + (i - 1) * e; AT(H, i, 0) = AT(I, i, 0); AT(D, i, 0) = NW_INF;
        for (int j = 1; j <= n; ++j) {
            int io = AT(H, i - 1, j) + o, ie = AT(I, i - 1, j) + e;
            int dop = AT(H, i, j - 1) + o, dex = AT(D, i, j - 1) + e;
            int iv = io < ie ? io : ie, dv = dop < dex ? dop : dex;
            int dg = AT(H, i - 1, j - 1) + (read[i - 1] != ref[j - 1] ? x : 0);
            int h = dg; if (iv < h) h = iv; if (dv < h) h = dv;
            AT(I, i, j) = iv; AT(D, i, j) = dv; AT(H, i, j) = h;
        }
    }
    int pen = AT(H, m, n);
    if (ops || nops_out) {
        /* traceback (m,n)->(0,0); collect reversed single ops, then RLE forwards */
        uint8_t* tr = (uint8_t*)malloc((size_t)m + (size_t)n + 2);
        int nt = 0, i = m, j = n, st = 0; /* st: 0=H, 1=I, 2=D */
        while (i > 0 || j > 0) {
            if (st == 0) {
                if (i > 0 && j > 0 && AT(H, i, j) == AT(H, i - 1, j - 1) + (read[i - 1] != ref[j - 1] ? x : 0)) {
                    tr[nt++] = (read[i - 1] != ref[j - 1]) ? GO_OP_X : GO_OP_EQ; --i; --j;
                } else if (i > 0 && AT(H, i, j) == AT(I, i, j)) st = 1;
                else st = 2;
            } else if (st == 1) {
                tr[nt++] = GO_OP_I;
                if (AT(I, i, j) == AT(H, i - 1, j) + o) st = 0; /* prefer closing the gap */
                --i;
            } else {
                tr[nt++] = GO_OP_D;
                if (AT(D, i, j) == AT(H, i, j - 1) + o) st = 0;
                --j;
            }
        }
        int nops = 0;
        for (int t = nt - 1; t >= 0;) {
            int code = tr[t], run = 0;
            while (t >= 0 && tr[t] == code) { ++run; --t; }
            if (ops && nops < cap) ops[nops] = (uint16_t)((run << 3) | code);
            ++nops;
        }
        if (nops_out) *nops_out = nops;
        free(tr);
    }
#undef AT
    free(H);
    return pen;
}

static void* nw_shard(void* arg) {
    shard_t* s = (shard_t*)arg;
    for (int64_t i = s->lo; i < s->hi; ++i) {
        int nops = 0;
        int pen = go_nw(s->reads + s->read_off[i], (int)(s->read_off[i + 1] - s->read_off[i]),
                        s->refs + s->ref_off[i], (int)(s->ref_off[i + 1] - s->ref_off[i]),
                        s->x, s->o, s->e, NULL, 0, s->op_cnt ? &nops : NULL);
        if (s->cost) s->cost[i] = pen;
        if (s->op_cnt) s->op_cnt[i] = (uint32_t)nops;
    }
    return NULL;
}

static void* nw_ops_shard(void* arg) {   /* op_cnt holds the CSR offsets here */
    shard_t* s = (shard_t*)arg;
    for (int64_t i = s->lo; i < s->hi; ++i) {
        int nops = 0;
        go_nw(s->reads + s->read_off[i], (int)(s->read_off[i + 1] - s->read_off[i]), s->refs + s->ref_off[i],
              (int)(s->ref_off[i + 1] - s->ref_off[i]), s->x, s->o, s->e, s->ops + s->op_cnt[i],
              (int)(s->op_cnt[i + 1] - s->op_cnt[i]), &nops);
    }
    return NULL;
}

int go_nw_batch(const char* reads, const uint32_t* read_off, const char* refs, const uint32_t* ref_off,
                int64_t n, int x, int o, int e, int threads, int32_t* penalty,
                uint16_t* ops, uint32_t* op_off, int64_t ops_cap) {
    shard_t proto; memset(&proto, 0, sizeof proto);
    proto.reads = reads; proto.read_off = read_off; proto.refs = refs; proto.ref_off = ref_off;
    proto.cost = penalty; proto.x = x; proto.o = o; proto.e = e;
    uint32_t* cnt = op_off ? (uint32_t*)malloc(sizeof(uint32_t) * (size_t)(n + 1)) : NULL;
    proto.op_cnt = cnt;
    shard_t sh[256];
    if (threads > 256) threads = 256;
    run_shards(nw_shard, &proto, n, threads, sh);
    if (op_off) {
        uint32_t acc = 0;
        for (int64_t i = 0; i < n; ++i) { op_off[i] = acc; acc += cnt[i]; }
        op_off[n] = acc;
        free(cnt);
        if (ops) {
            if ((int64_t)acc > ops_cap) return -3;
            /* second pass, sharded like the first: every pair writes its ops at its CSR offset */
            proto.ops = ops; proto.op_cnt = op_off; proto.cost = NULL;
            run_shards(nw_ops_shard, &proto, n, threads, sh);
        }
    }
    return 0;
}

/* ------------------------------------------------------------------------------------------
 * LEAP comparator: banded affine Landau-Vishkin, class LV of GASMA/benchmark/LEAP_SIMD/LV_BAG.cpp,
 * driven as benchmark::_run_LEAP does (benchmark_utils.h:156-178) on an object set up with
 * init(k, af_threshold = 200, ED_GLOBAL, x, o, e) (benchmark_utils.h:289).
 *   lanes: 2k+3 of them, lane k+1 is the main diagonal; lanes 0 and 2k+2 are never-valid sentinels
 *   (LV_BAG.cpp:76-77,129).  Below the middle the read is offset, above it the reference (:9-23).
 *   load_reads pads both strings with NUL up to L = max(m, n) (strncpy, :112-123), so a position past
 *   the end of the shorter string mismatches.
 *   cells start/end/I_pos/D_pos[lane][e] are -2 until reached; which cells get reached depends only on
 *   (lane, e, penalties), never on the data, so reusing the object across pairs (as the benchmark does)
 *   gives the same results as a fresh one -- checked against the reference in tests/test_oracle.py.
 * Returns the penalty the benchmark records (get_ED() = final_ED, :360: the error level at which the
 * best converging lane reached the end, WITHOUT the cost of returning to the main lane), or INT32_MIN
 * when no lane converges within af_threshold (the reference would then report the previous pair's
 * value).  Needs o >= e ("Assuming gap_open_penalty > gap_ext_penalty", :165) and m, n <= 256.
 * ---------------------------------------------------------------------------------------- */
#define LV_MAXLEN 256
#define LV_AF 200
int go_leap(const char* read, int m, const char* ref, int n, int k, int x, int o, int e) {
    if (m > LV_MAXLEN || n > LV_MAXLEN || k < 0 || k > GO_MAX_K || x < 1 || e < 1 || o < e) return INT32_MIN;
    const int total = 2 * k + 3, mid = k + 1, L = m > n ? m : n;
    char A[LV_MAXLEN + 1], B[LV_MAXLEN + 1];
    memset(A, 0, sizeof A); memset(B, 0, sizeof B);
    memcpy(A, read, (size_t)m); memcpy(B, ref, (size_t)n);
    /* [lane][e] tables, -2 = not reached */
    static __thread int st[2 * GO_MAX_K + 3][LV_AF + 1], en[2 * GO_MAX_K + 3][LV_AF + 1];
    static __thread int ip[2 * GO_MAX_K + 3][LV_AF + 1], dp[2 * GO_MAX_K + 3][LV_AF + 1];
    for (int l = 0; l < total; ++l)
        for (int q = 0; q <= LV_AF; ++q) st[l][q] = en[l][q] = ip[l][q] = dp[l][q] = -2;
    st[mid][0] = 0;   /* ED_GLOBAL: only the main lane starts at error level 0 (:99-101) */
#define LV_EXT(l, pos, out) do { \
        int ao_ = (l) < mid ? mid - (l) : 0, bo_ = (l) > mid ? (l) - mid : 0, p_ = (pos); \
        while (p_ < L && p_ - ao_ >= 0 && p_ - bo_ >= 0 && A[p_ - ao_] == B[p_ - bo_]) ++p_; \
        (out) = p_; } while (0)
    int pass = 0, final_ed = 0, converge = 1000000;
    for (int l = 1; l < total - 1; ++l) {
        if (st[l][0] >= 0) {
            LV_EXT(l, st[l][0], en[l][0]);
            if (en[l][0] == L) return 0;
        }
    }
    for (int q = 1; q <= LV_AF; ++q) {
        for (int l = 1; l < total - 1; ++l) {
            const int top = l >= mid, bot = l <= mid;
            if (q >= o && en[l - 1][q - o] >= 0 && en[l - 1][q - o] > ip[l - 1][q - e]) ip[l][q] = en[l - 1][q - o] + top;
            else if (q >= e && ip[l - 1][q - e] >= 0) ip[l][q] = ip[l - 1][q - e] + top;
            if (q >= o && en[l + 1][q - o] >= 0 && en[l + 1][q - o] > dp[l + 1][q - e]) dp[l][q] = en[l + 1][q - o] + bot;
            else if (q >= e && dp[l + 1][q - e] >= 0) dp[l][q] = dp[l + 1][q - e] + bot;
            int s0 = -2;
            if (q >= x && en[l][q - x] >= 0) s0 = en[l][q - x] + 1;
            if (ip[l][q] > s0) s0 = ip[l][q];
            if (dp[l][q] > s0) s0 = dp[l][q];
            st[l][q] = s0;
            if (s0 >= 0) {
                LV_EXT(l, s0, en[l][q]);
                if (en[l][q] == L) {
                    const int d = abs(mid - l);
                    const int conv = q + (d ? o + (d - 1) * e : 0);
                    if (conv <= LV_AF && conv < converge) { final_ed = q; pass = 1; converge = conv; }
                }
            }
        }
        if (pass) break;
    }
#undef LV_EXT
    return pass ? final_ed : INT32_MIN;
}

static void* leap_shard(void* arg) {
    shard_t* s = (shard_t*)arg;
    for (int64_t i = s->lo; i < s->hi; ++i)
        s->cost[i] = go_leap(s->reads + s->read_off[i], (int)(s->read_off[i + 1] - s->read_off[i]),
                             s->refs + s->ref_off[i], (int)(s->ref_off[i + 1] - s->ref_off[i]), s->stride, s->x, s->o, s->e);
    return NULL;
}

int go_leap_batch(const char* reads, const uint32_t* read_off, const char* refs, const uint32_t* ref_off,
                  int64_t n, int k, int x, int o, int e, int threads, int32_t* penalty) {
    shard_t proto; memset(&proto, 0, sizeof proto);
    proto.reads = reads; proto.read_off = read_off; proto.refs = refs; proto.ref_off = ref_off;
    proto.cost = penalty; proto.x = x; proto.o = o; proto.e = e; proto.stride = k;
    shard_t sh[256];
    if (threads > 256) threads = 256;
    run_shards(leap_shard, &proto, n, threads, sh);
    return 0;
}

/* ------------------------------------------------------------------------------------------
 * coverage metric, benchmark_coverage.h:26-67 and :73-91, on encoded ops
 * ---------------------------------------------------------------------------------------- */
int go_lcm(const char* s1, int s1_len, const uint16_t* ops, int nops, int threshold, char* out, int cap) {
    int s1_index = 0, pos = 0;
    for (int t = 0; t < nops; ++t) {
        int length = ops[t] >> 3, code = ops[t] & 7;
        switch (code) {
            case GO_OP_X: s1_index += length; break;
            case GO_OP_I: s1_index += length; break;
            case GO_OP_D: break;
            default: /* '=' or 'M' */
                for (int i = 0; i < length; ++i) {
                    if (length >= threshold) {
                        /* the reference indexes s1 unchecked; reads past s1_len yield NUL from std::string */
                        char c = s1_index < s1_len ? s1[s1_index] : '\0';
                        if (pos < cap) out[pos] = c;
                        ++pos;
                    }
                    ++s1_index;
                }
        }
    }
    return pos;
}

int go_covers(const char* a, int na, const char* b, int nb) {
    if (na < nb) return 0;
    int i = 0;
    for (int j = 0; j < nb; ++j) {
        if (i >= na) return 0;
        while (a[i] != b[j]) { ++i; if (i >= na) return 0; }
        ++i;
    }
    return 1;
}

/* benchmark_utils.h:231-259 */
void go_metrics(const char* reads, const uint32_t* read_off, int64_t n,
                const int32_t* greedy_cost, const uint16_t* g_ops, const uint32_t* g_off,
                const int32_t* nw_penalty, const uint16_t* n_ops, const uint32_t* n_off,
                const int32_t* answers, uint64_t out[4]) {
    out[0] = out[1] = out[2] = out[3] = 0;
    char* l1 = (char*)malloc(64 * GO_MAX_W * 2 + 16); char* l2 = (char*)malloc(64 * GO_MAX_W * 2 + 16);
    for (int64_t i = 0; i < n; ++i) {
        int correct = (answers && answers[i] != INT32_MIN) ? answers[i] : nw_penalty[i];
        out[0] += 1;
        out[1] += (nw_penalty[i] == correct);
        out[2] += (greedy_cost[i] == correct);
        if (g_ops && n_ops) {
            const char* s1 = reads + read_off[i]; int s1_len = (int)(read_off[i + 1] - read_off[i]);
            int a = go_lcm(s1, s1_len, g_ops + g_off[i], (int)(g_off[i + 1] - g_off[i]), 1, l1, 64 * GO_MAX_W * 2);
            int b = go_lcm(s1, s1_len, n_ops + n_off[i], (int)(n_off[i + 1] - n_off[i]), 3, l2, 64 * GO_MAX_W * 2);
            out[3] += (uint64_t)go_covers(l1, a, l2, b);
        }
    }
    free(l1); free(l2);
}

/* ------------------------------------------------------------------------------------------
 * Synthetic pairs.  Restates Dataset (benchmark_dataset.h:100-186): read = L iid uniform ACGT;
 * ref = copy of the read, then edits applied sequentially to the ref: with probability
 * mismatch_rate overwrite a random position with a random base (may equal the old base, :113-120),
 * otherwise delete (:121-132) or insert (:133-146) at a random position, 50/50 (:169).
 * libc rand() is replaced by a counter RNG (splitmix64 stream keyed by seed and pair index) so
 * that any shard regenerates pair i identically on CPU and GPU.  Mode 2 is a per-base
 * independent error profile (README.md:71-76, SRR611076-shaped).
 * The device twin is approximate-string-matching_b200/csrc/generate.cuh; they must stay in lockstep.
 * ---------------------------------------------------------------------------------------- */
typedef struct { uint64_t s; } rng_t;
static uint64_t rng_next(rng_t* r) {
    uint64_t z = (r->s += 0x9E3779B97F4A7C15ull);
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}
static uint32_t rng_u32(rng_t* r) { return (uint32_t)(rng_next(r) >> 32); }
static uint32_t rng_below(rng_t* r, uint32_t n) { return (uint32_t)(((uint64_t)rng_u32(r) * n) >> 32); }
static uint64_t rng_mix(uint64_t z) {
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}
/* a pair's stream starts at a hashed point of the cycle, so that the streams of neighbouring pairs do not overlap
 * (twin of Rng::seed in csrc/generate.cuh) */
static void rng_seed(rng_t* r, uint64_t seed, uint64_t pair) {
    r->s = rng_mix(rng_mix(seed + 0x2545F4914F6CDD1Dull) ^ (pair * 0xD1B54A32D192ED03ull + 0x9E3779B97F4A7C15ull));
}

int go_gen_ref_capacity(const go_gen_config* c) {
    return c->mode == 2 ? 2 * c->length + 8 : c->length + c->num_errors + 1;
}

int go_generate_pair(const go_gen_config* c, uint64_t pair_index, char* read, char* ref) {
    static const char alphabet[4] = {'A', 'C', 'G', 'T'};
    rng_t r; rng_seed(&r, c->seed, pair_index);
    const int L = c->length;
    for (int i = 0; i < L; ++i) read[i] = alphabet[rng_u32(&r) >> 30];
    int len;
    if (c->mode == 2) {
        len = 0;
        for (int i = 0; i < L; ++i) {
            uint64_t u = rng_u32(&r);
            if (u < c->thr_del) continue;                                 /* base missing from the ref */
            if (u < c->thr_ins) { ref[len++] = alphabet[rng_u32(&r) >> 30]; ref[len++] = read[i]; continue; }
            if (u < c->thr_sub) {                                         /* one of the 3 other bases */
                int b = (read[i] == 'A') ? 0 : (read[i] == 'C') ? 1 : (read[i] == 'G') ? 2 : 3;
                ref[len++] = alphabet[(b + 1 + (int)rng_below(&r, 3)) & 3];
                continue;
            }
            ref[len++] = read[i];
        }
        return len;
    }
    memcpy(ref, read, (size_t)L);
    len = L;
    int ne = c->num_errors;
    if (c->mode == 1) ne = c->num_errors > 0 ? (int)rng_below(&r, (uint32_t)c->num_errors) : 0;
    for (int t = 0; t < ne; ++t) {
        uint64_t u = rng_u32(&r);
        if (u < c->thr_mismatch) {
            if (len == 0) continue;
            int pos = (int)rng_below(&r, (uint32_t)len);
            ref[pos] = alphabet[rng_u32(&r) >> 30];
        } else if (rng_below(&r, 2) == 0) { /* deletion */
            if (len == 0) continue;
            int pos = (int)rng_below(&r, (uint32_t)len);
            memmove(ref + pos, ref + pos + 1, (size_t)(len - 1 - pos));
            --len;
        } else {                            /* insertion */
            int pos = len ? (int)rng_below(&r, (uint32_t)len) : 0;
            memmove(ref + pos + 1, ref + pos, (size_t)(len - pos));
            ++len;
            ref[pos] = alphabet[rng_u32(&r) >> 30];
        }
    }
    return len;
}
