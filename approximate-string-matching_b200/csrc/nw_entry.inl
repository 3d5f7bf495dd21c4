// nw_entry.inl -- C-ABI entry points of K4 (NW verifier) and K5 (counters); included by gasma_b200.cu.
namespace {

template <int W>
int launch_nw_w(gasma_ctx* ctx, const uint8_t* reads, const uint32_t* roff, const uint8_t* refs, const uint32_t* foff,
                int64_t n, const int32_t* bound, int32_t* penalty, uint16_t* stage, int stride, uint32_t* nops) {
    constexpr int T = 128;
    constexpr int BITS = 64 * W;
    uint32_t* misc = (uint32_t*)ctx->misc.p;
    uint32_t* dcount = misc + 4;
    const bool want_ops = stage != nullptr;
    int rc;
    if ((rc = reserve(ctx, ctx->nw_deferred, (size_t)n * 4))) return rc;
    const int nslots = W <= 2 ? 8192 : 4096;
    if ((rc = reserve(ctx, ctx->nw_pool_hi, (size_t)nslots * 2 * (2 * BITS + 2) * 4))) return rc;
    if (want_ops && (rc = reserve(ctx, ctx->nw_pool_tb, (size_t)nslots * BITS * nw_tb_words_wide<W>() * 4))) return rc;
    if (ctx->kp.x == 1 && ctx->kp.o == 1 && ctx->kp.e == 1) {
        // unit penalties (the benchmark's configuration): bit-vector columns.
        //   with a bound (the greedy penalty): one banded pass gives penalty and traceback;
        //   without: a penalty-only pass over full columns first, whose exact optimum then is the bound.
        // Pairs whose band exceeds the 32-bit window (rare) finish in the general Gotoh kernel.
        cudaEvent_t* mev = ctx->ev[1][ctx->ev_count[1] % gasma_ctx::kRing];
        CU(cudaEventRecord(mev[0], ctx->stream));
        if (!bound) {
            auto mk = nw_myers_kernel<W, T>;
            const size_t msmem = myers_smem_bytes<W>(T);
            CU(cudaFuncSetAttribute(mk, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)msmem));
            mk<<<(unsigned)((n + T - 1) / T), T, msmem, ctx->stream>>>(reads, roff, refs, foff, (long long)n, penalty, misc);
            CU(cudaGetLastError());
            ctx->launches++;
            bound = penalty;
        }
        if (want_ops || bound != penalty) {
            CU(cudaMemsetAsync(dcount, 0, 4, ctx->stream));
            if (want_ops) {
                constexpr int TT = 64;
                auto tk = nw_band_kernel<W, TT, kNwOps>;
                const size_t tsmem = myers_trace_smem_bytes<W>(TT);
                CU(cudaFuncSetAttribute(tk, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)tsmem));
                tk<<<(unsigned)((n + TT - 1) / TT), TT, tsmem, ctx->stream>>>(reads, roff, refs, foff, (long long)n, bound, penalty, stage,
                                                                            stride, nops, (uint32_t*)ctx->nw_deferred.p, dcount, misc, NwCounterArgs{});
            } else {
                auto tk = nw_band_kernel<W, T, kNwPenalty>;
                const size_t tsmem = myers_smem_bytes<W>(T);
                CU(cudaFuncSetAttribute(tk, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)tsmem));
                tk<<<(unsigned)((n + T - 1) / T), T, tsmem, ctx->stream>>>(reads, roff, refs, foff, (long long)n, bound, penalty, nullptr,
                                                                          0, nullptr, (uint32_t*)ctx->nw_deferred.p, dcount, misc, NwCounterArgs{});
            }
            CU(cudaGetLastError());
            nw_wide_kernel<W><<<nslots / 128, 128, 0, ctx->stream>>>(
                1, 1, 1, reads, roff, refs, foff, penalty, penalty, (int32_t*)ctx->nw_pool_hi.p, (uint32_t*)ctx->nw_pool_tb.p, nslots,
                stage, stride, nops, (const uint32_t*)ctx->nw_deferred.p, dcount);
            CU(cudaGetLastError());
            ctx->launches += 2;
        }
        CU(cudaEventRecord(mev[1], ctx->stream));
        ctx->ev_count[1]++;
        return GASMA_OK;
    }
    // traceback scratch of the banded Gotoh kernel: bounded to ~1 GiB by processing the batch in chunks
    const int64_t chunk = want_ops ? ((int64_t)1 << 30) / ((int64_t)BITS * nw_tb_words_fast<W>() * 4) : n;
    if (want_ops && (rc = reserve(ctx, ctx->nw_tb, (size_t)(n < chunk ? n : chunk) * BITS * nw_tb_words_fast<W>() * 4))) return rc;
    CU(cudaMemsetAsync(dcount, 0, 4, ctx->stream));
    auto kern = nw_kernel<W, T>;
    const size_t smem = nw_smem_bytes<W>(T);
    CU(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    cudaEvent_t* evp = ctx->ev[1][ctx->ev_count[1] % gasma_ctx::kRing];
    CU(cudaEventRecord(evp[0], ctx->stream));
    for (int64_t first = 0; first < n; first += chunk) {
        const int64_t count = n - first < chunk ? n - first : chunk;
        kern<<<(unsigned)((count + T - 1) / T), T, smem, ctx->stream>>>(
            ctx->kp.x, ctx->kp.o, ctx->kp.e, reads, roff, refs, foff, (long long)n, (long long)first, (long long)count, bound,
            penalty, (uint32_t*)ctx->nw_tb.p, stage, stride, nops, (uint32_t*)ctx->nw_deferred.p, dcount, misc);
        CU(cudaGetLastError());
        ctx->launches++;
    }
    nw_wide_kernel<W><<<nslots / 128, 128, 0, ctx->stream>>>(
        ctx->kp.x, ctx->kp.o, ctx->kp.e, reads, roff, refs, foff, bound, penalty, (int32_t*)ctx->nw_pool_hi.p,
        (uint32_t*)ctx->nw_pool_tb.p, nslots, stage, stride, nops, (const uint32_t*)ctx->nw_deferred.p, dcount);
    CU(cudaGetLastError());
    ctx->launches++;
    CU(cudaEventRecord(evp[1], ctx->stream));
    ctx->ev_count[1]++;
    return GASMA_OK;
}

// The benchmark body's verifier (unit penalties): NW penalties + the four counters in one kernel, bounded by the
// greedy penalties; the NW CIGAR only exists for the few pairs the banded kernel defers.
template <int W>
int launch_nw_counters_w(gasma_ctx* ctx, const uint8_t* reads, const uint32_t* roff, const uint8_t* refs, const uint32_t* foff,
                         int64_t n, const int32_t* gcost, int32_t* penalty, const NwCounterArgs& ca, uint16_t* stage, int stride,
                         uint32_t* nops) {
    constexpr int TT = 64;
    constexpr int BITS = 64 * W;
    uint32_t* misc = (uint32_t*)ctx->misc.p;
    uint32_t* dcount = misc + 4;
    int rc;
    if ((rc = reserve(ctx, ctx->nw_deferred, (size_t)n * 4))) return rc;
    const int nslots = W <= 2 ? 8192 : 4096;
    if ((rc = reserve(ctx, ctx->nw_pool_hi, (size_t)nslots * 2 * (2 * BITS + 2) * 4))) return rc;
    if ((rc = reserve(ctx, ctx->nw_pool_tb, (size_t)nslots * BITS * nw_tb_words_wide<W>() * 4))) return rc;
    cudaEvent_t* mev = ctx->ev[1][ctx->ev_count[1] % gasma_ctx::kRing];
    CU(cudaEventRecord(mev[0], ctx->stream));
    CU(cudaMemsetAsync(dcount, 0, 8, ctx->stream));   // deferred count + hard count (misc[4], misc[5])
    auto tk = nw_band_kernel<W, TT, kNwCounters>;
    const size_t tsmem = myers_trace_smem_bytes<W>(TT);
    CU(cudaFuncSetAttribute(tk, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)tsmem));
    tk<<<(unsigned)((n + TT - 1) / TT), TT, tsmem, ctx->stream>>>(reads, roff, refs, foff, (long long)n, gcost, penalty, nullptr, stride,
                                                                nullptr, (uint32_t*)ctx->nw_deferred.p, dcount, misc, ca);
    CU(cudaGetLastError());
    nw_wide_kernel<W><<<nslots / 128, 128, 0, ctx->stream>>>(
        1, 1, 1, reads, roff, refs, foff, penalty, penalty, (int32_t*)ctx->nw_pool_hi.p, (uint32_t*)ctx->nw_pool_tb.p, nslots,
        stage, stride, nops, (const uint32_t*)ctx->nw_deferred.p, dcount);
    CU(cudaGetLastError());
    nw_deferred_counters_kernel<<<8, 128, 0, ctx->stream>>>(reads, roff, gcost, penalty, stage, stride, nops, ca,
                                                           (const uint32_t*)ctx->nw_deferred.p, dcount);
    CU(cudaGetLastError());
    // sized for the worst case (every pair queued); blocks beyond the queue's length return at once
    cover_hard_kernel<W><<<(unsigned)((n + 127) / 128), 128, 0, ctx->stream>>>(ca.hard, ca.hard_count, ca.hard_cap, ca.counters);
    CU(cudaGetLastError());
    ctx->launches += 4;
    CU(cudaEventRecord(mev[1], ctx->stream));
    ctx->ev_count[1]++;
    return GASMA_OK;
}

// The greedy CIGAR is read from the staging rows the greedy step of the same context just filled (rows_only mode of
// greedy_dev_impl): ctx->ops_rows / ops_stage / nops.  The verifier's own rows for deferred pairs go to bm_nstage.
int nw_counters_dev_impl(gasma_ctx* ctx, const char* d_reads, const uint32_t* d_read_off, const char* d_refs,
                         const uint32_t* d_ref_off, int64_t n, const int32_t* d_gcost, int32_t* d_penalty,
                         const int32_t* d_answers, uint64_t* d_counters) {
    const int stride = 2 * ctx->params.max_length + 8;
    int rc;
    if ((rc = reserve(ctx, ctx->bm_nstage, (size_t)n * stride * sizeof(uint16_t)))) return rc;
    if ((rc = reserve(ctx, ctx->bm_nnops, (size_t)n * 4))) return rc;
    if ((rc = reserve(ctx, ctx->bm_hard, (size_t)n * (8 * ctx->W + 1) * 4))) return rc;
    const NwCounterArgs ca{(const uint16_t*)ctx->ops_rows.p, (const uint16_t*)ctx->ops_stage.p, stride, (const uint32_t*)ctx->nops.p,
                           d_answers, (unsigned long long*)d_counters,
                           (uint32_t*)ctx->bm_hard.p, (uint32_t*)ctx->misc.p + 5, (uint32_t)n, (uint32_t*)ctx->misc.p + 6};
    const uint8_t* r = (const uint8_t*)d_reads;
    const uint8_t* f = (const uint8_t*)d_refs;
    uint16_t* stage = (uint16_t*)ctx->bm_nstage.p;
    uint32_t* nops = (uint32_t*)ctx->bm_nnops.p;
    switch (ctx->W) {
        case 1: rc = launch_nw_counters_w<1>(ctx, r, d_read_off, f, d_ref_off, n, d_gcost, d_penalty, ca, stage, stride, nops); break;
        case 2: rc = launch_nw_counters_w<2>(ctx, r, d_read_off, f, d_ref_off, n, d_gcost, d_penalty, ca, stage, stride, nops); break;
        case 3: rc = launch_nw_counters_w<3>(ctx, r, d_read_off, f, d_ref_off, n, d_gcost, d_penalty, ca, stage, stride, nops); break;
        case 4: rc = launch_nw_counters_w<4>(ctx, r, d_read_off, f, d_ref_off, n, d_gcost, d_penalty, ca, stage, stride, nops); break;
        default: return fail(ctx, GASMA_E_UNSUPPORTED, "the NW verifier covers max_length <= 256 (the greedy path covers 512 and 1024)");
    }
    if (rc == GASMA_OK) ctx->counter_clean = true;   // the verifier reset the greedy kernel's chunk counter
    return rc;
}

int nw_dev_impl(gasma_ctx* ctx, const char* d_reads, const uint32_t* d_read_off, const char* d_refs,
                const uint32_t* d_ref_off, int64_t n, const int32_t* d_bound, int32_t* d_penalty,
                gasma_cigar_op* d_ops, int64_t ops_cap, uint32_t* d_op_off) {
    if (!ctx) return GASMA_E_ARG;
    if (n < 0 || n > (int64_t)1 << 31) return fail(ctx, GASMA_E_ARG, "n_pairs out of range");
    if (!d_read_off || !d_ref_off || !d_penalty) return fail(ctx, GASMA_E_ARG, "null offsets/penalty pointer");
    if ((d_ops != nullptr) != (d_op_off != nullptr)) return fail(ctx, GASMA_E_ARG, "ops and op_off must be given together");
    if (!aligned16(d_reads) || !aligned16(d_refs)) return fail(ctx, GASMA_E_ARG, "sequence buffers must be 16-byte aligned");
    CU(cudaSetDevice(ctx->device));
    if (n == 0) {
        if (d_op_off) CU(cudaMemsetAsync(d_op_off, 0, 4, ctx->stream));
        return GASMA_OK;
    }
    if (ctx->W > 4)
        return fail(ctx, GASMA_E_UNSUPPORTED, "the NW verifier covers max_length <= 256 (the greedy path covers 512 and 1024)");
    const int T = 128;
    const int64_t ntiles = (n + T - 1) / T;
    const int stride = 2 * ctx->params.max_length + 8;
    const bool want_ops = d_ops != nullptr;
    int rc;
    if (want_ops) {
        if ((rc = reserve(ctx, ctx->ops_stage, (size_t)n * stride * sizeof(uint16_t)))) return rc;
        if ((rc = reserve(ctx, ctx->nops, (size_t)n * 4))) return rc;
        if ((rc = reserve(ctx, ctx->tile_sums, (size_t)ntiles * 4))) return rc;
        if ((rc = reserve(ctx, ctx->tile_bases, (size_t)ntiles * 4))) return rc;
    }
    uint16_t* stage = want_ops ? (uint16_t*)ctx->ops_stage.p : nullptr;
    uint32_t* nops = want_ops ? (uint32_t*)ctx->nops.p : nullptr;
    const uint8_t* r = (const uint8_t*)d_reads;
    const uint8_t* f = (const uint8_t*)d_refs;
    switch (ctx->W) {
        case 1: rc = launch_nw_w<1>(ctx, r, d_read_off, f, d_ref_off, n, d_bound, d_penalty, stage, stride, nops); break;
        case 2: rc = launch_nw_w<2>(ctx, r, d_read_off, f, d_ref_off, n, d_bound, d_penalty, stage, stride, nops); break;
        case 3: rc = launch_nw_w<3>(ctx, r, d_read_off, f, d_ref_off, n, d_bound, d_penalty, stage, stride, nops); break;
        default: rc = launch_nw_w<4>(ctx, r, d_read_off, f, d_ref_off, n, d_bound, d_penalty, stage, stride, nops); break;
    }
    if (rc) return rc;
    if (want_ops) {
        uint32_t* misc = (uint32_t*)ctx->misc.p;
        tile_sums_kernel<128><<<(unsigned)ntiles, 128, 0, ctx->stream>>>(nops, (long long)n, (uint32_t*)ctx->tile_sums.p);
        CU(cudaGetLastError());
        scan_tiles_kernel<<<1, 1024, 0, ctx->stream>>>((uint32_t*)ctx->tile_sums.p, (uint32_t*)ctx->tile_bases.p, (int)ntiles, misc + 1, nullptr);
        CU(cudaGetLastError());
        compact_ops_kernel<128, true><<<(unsigned)ntiles, 128, 0, ctx->stream>>>(stage, stride, nops, (uint32_t*)ctx->tile_bases.p,
                                                                         (long long)n, d_ops, (long long)ops_cap, d_op_off, 1);
        CU(cudaGetLastError());
        ctx->launches += 3;
    }
    return GASMA_OK;
}

}  // namespace

extern "C" {

int gasma_nw_batch_dev(gasma_ctx* ctx, const char* d_reads, const uint32_t* d_read_off, const char* d_refs,
                       const uint32_t* d_ref_off, int64_t n_pairs, const int32_t* d_bound, int32_t* d_penalty,
                       gasma_cigar_op* d_ops, int64_t ops_cap, uint32_t* d_op_off) {
    return nw_dev_impl(ctx, d_reads, d_read_off, d_refs, d_ref_off, n_pairs, d_bound, d_penalty, d_ops, ops_cap, d_op_off);
}

int gasma_nw_batch(gasma_ctx* ctx, const char* reads, const uint32_t* read_off, const char* refs, const uint32_t* ref_off,
                   int64_t n, const int32_t* bound, int32_t* penalty, gasma_cigar_op* ops, int64_t ops_cap, uint32_t* op_off) {
    if (!ctx) return GASMA_E_ARG;
    if (n < 0 || !read_off || !ref_off || !penalty) return fail(ctx, GASMA_E_ARG, "bad argument");
    if ((ops != nullptr) != (op_off != nullptr)) return fail(ctx, GASMA_E_ARG, "ops and op_off must be given together");
    if (n == 0) { if (op_off) op_off[0] = 0; return GASMA_OK; }
    int rc;
    if ((rc = check_offsets(ctx, read_off, n, "read"))) return rc;
    if ((rc = check_offsets(ctx, ref_off, n, "ref"))) return rc;
    CU(cudaSetDevice(ctx->device));
    const size_t rb = read_off[n], fb = ref_off[n], ob = (size_t)(n + 1) * 4;
    if ((rc = reserve(ctx, ctx->d_reads, rb + 16))) return rc;
    if ((rc = reserve(ctx, ctx->d_refs, fb + 16))) return rc;
    if ((rc = reserve(ctx, ctx->d_read_off, ob))) return rc;
    if ((rc = reserve(ctx, ctx->d_ref_off, ob))) return rc;
    if ((rc = reserve(ctx, ctx->d_cost, (size_t)n * 4))) return rc;
    if (bound && (rc = reserve(ctx, ctx->d_bound, (size_t)n * 4))) return rc;
    if (ops) {
        if ((rc = reserve(ctx, ctx->d_op_off, ob))) return rc;
        if ((rc = reserve(ctx, ctx->d_ops, (size_t)(ops_cap > 0 ? ops_cap : 1) * 2))) return rc;
    }
    if (rb) CU(cudaMemcpyAsync(ctx->d_reads.p, reads, rb, cudaMemcpyHostToDevice, ctx->stream));
    if (fb) CU(cudaMemcpyAsync(ctx->d_refs.p, refs, fb, cudaMemcpyHostToDevice, ctx->stream));
    CU(cudaMemcpyAsync(ctx->d_read_off.p, read_off, ob, cudaMemcpyHostToDevice, ctx->stream));
    CU(cudaMemcpyAsync(ctx->d_ref_off.p, ref_off, ob, cudaMemcpyHostToDevice, ctx->stream));
    if (bound) CU(cudaMemcpyAsync(ctx->d_bound.p, bound, (size_t)n * 4, cudaMemcpyHostToDevice, ctx->stream));
    rc = nw_dev_impl(ctx, (const char*)ctx->d_reads.p, (const uint32_t*)ctx->d_read_off.p, (const char*)ctx->d_refs.p,
                     (const uint32_t*)ctx->d_ref_off.p, n, bound ? (const int32_t*)ctx->d_bound.p : nullptr,
                     (int32_t*)ctx->d_cost.p, ops ? (gasma_cigar_op*)ctx->d_ops.p : nullptr, ops_cap,
                     ops ? (uint32_t*)ctx->d_op_off.p : nullptr);
    if (rc) return rc;
    CU(cudaMemcpyAsync(penalty, ctx->d_cost.p, (size_t)n * 4, cudaMemcpyDeviceToHost, ctx->stream));
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

int gasma_metrics_dev(gasma_ctx* ctx, const char* d_reads, const uint32_t* d_read_off, int64_t n,
                      const int32_t* d_greedy_cost, const gasma_cigar_op* d_greedy_ops, const uint32_t* d_greedy_off,
                      const int32_t* d_nw_penalty, const gasma_cigar_op* d_nw_ops, const uint32_t* d_nw_off,
                      const int32_t* d_answers, uint64_t* d_counters) {
    if (!ctx) return GASMA_E_ARG;
    if (n < 0 || !d_greedy_cost || !d_nw_penalty || !d_counters) return fail(ctx, GASMA_E_ARG, "bad argument");
    const bool cov = d_greedy_ops && d_nw_ops;
    if (cov && (!d_reads || !d_read_off || !d_greedy_off || !d_nw_off)) return fail(ctx, GASMA_E_ARG, "coverage needs reads and both CSR arrays");
    if (cov && !aligned16(d_reads)) return fail(ctx, GASMA_E_ARG, "sequence buffers must be 16-byte aligned");
    CU(cudaSetDevice(ctx->device));
    if (n == 0) return GASMA_OK;
    if (cov && ctx->W > 4) return fail(ctx, GASMA_E_UNSUPPORTED, "coverage counters cover max_length <= 256");
    const int smem_cap = (128 * ctx->params.max_length + 32 + 15) / 16 * 16;   // a tile's reads (+ alignment slack)
    const int smem_all = smem_cap + 2 * kCovWords * 128 * 4;                   // + the position masks
    CU(cudaFuncSetAttribute(metrics_kernel<128>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_all));
    metrics_kernel<128><<<(unsigned)((n + 127) / 128), 128, smem_all, ctx->stream>>>(
        (const uint8_t*)d_reads, d_read_off, (long long)n, d_greedy_cost, cov ? d_greedy_ops : nullptr, d_greedy_off,
        d_nw_penalty, cov ? d_nw_ops : nullptr, d_nw_off, d_answers, (unsigned long long*)d_counters, smem_cap);
    CU(cudaGetLastError());
    ctx->launches++;
    return GASMA_OK;
}

}  // extern "C"
