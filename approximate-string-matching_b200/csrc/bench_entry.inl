// bench_entry.inl -- C-ABI entry points of the batched benchmark body (greedy + NW verifier + counters in one
// call) and the host-pointer generator; included by gasma_b200.cu.
namespace {

// device time of the most recent recorded launch group of `which` (0 greedy, 1 NW); the stream must be idle
double last_kernel_ms(gasma_ctx* ctx, int which) {
    if (ctx->ev_count[which] <= 0) return 0.0;
    cudaEvent_t* e = ctx->ev[which][(ctx->ev_count[which] - 1) % gasma_ctx::kRing];
    float ms = 0.f;
    if (cudaEventElapsedTime(&ms, e[0], e[1]) != cudaSuccess) return 0.0;
    return ms;
}

// greedy -> NW (band from the greedy penalty) -> counters, all enqueued on the context's stream without a host
// round trip.  With unit penalties the verifier kernel also produces the counters (no NW CIGAR is materialised);
// other penalties run verifier and counter kernel separately.  The CIGAR CSR arrays are sized for the worst case (2*max_length+8 ops per pair, the staging row
// of either kernel), so no batch can overflow them and no op total has to travel to the host before the counters run.
int benchmark_dev_impl(gasma_ctx* ctx, const char* d_reads, const uint32_t* d_read_off, const char* d_refs,
                       const uint32_t* d_ref_off, int64_t n, const int32_t* d_answers, uint64_t* d_counters,
                       int32_t* d_gcost, int32_t* d_nwpen) {
    if (!ctx) return GASMA_E_ARG;
    if (n < 0 || !d_read_off || !d_ref_off || !d_counters) return fail(ctx, GASMA_E_ARG, "bad argument");
    if (ctx->kp.o < 1 || ctx->kp.e < 1)
        return fail(ctx, GASMA_E_UNSUPPORTED, "the banded verifier needs o,e >= 1 (use gasma_nw_batch with bound=NULL)");
    if (n == 0) return GASMA_OK;
    CU(cudaSetDevice(ctx->device));
    int rc;
    const size_t ob = (size_t)(n + 1) * 4;
    if (!d_gcost) { if ((rc = reserve(ctx, ctx->bm_gcost, (size_t)n * 4))) return rc; d_gcost = (int32_t*)ctx->bm_gcost.p; }
    if (!d_nwpen) { if ((rc = reserve(ctx, ctx->bm_nwpen, (size_t)n * 4))) return rc; d_nwpen = (int32_t*)ctx->bm_nwpen.p; }
    if (ctx->kp.x == 1 && ctx->kp.o == 1 && ctx->kp.e == 1 && ctx->W <= 4) {
        // unit penalties (what the reference's benchmark always runs): verifier and counters are one kernel, and it reads
        // the greedy CIGAR straight from the staging rows -- no CSR is built for it (two dominant launches per batch)
        if ((rc = greedy_dev_impl(ctx, d_reads, d_read_off, d_refs, d_ref_off, n, d_gcost, nullptr, 0, nullptr, false, /*rows_only=*/true))) return rc;
        return nw_counters_dev_impl(ctx, d_reads, d_read_off, d_refs, d_ref_off, n, d_gcost, d_nwpen, d_answers, d_counters);
    }
    if ((rc = reserve(ctx, ctx->bm_goff, ob))) return rc;
    const int64_t cap = n * (2 * (int64_t)ctx->params.max_length + 8);
    if ((rc = reserve(ctx, ctx->bm_gops, (size_t)cap * 2))) return rc;
    uint32_t* goff = (uint32_t*)ctx->bm_goff.p;
    if ((rc = greedy_dev_impl(ctx, d_reads, d_read_off, d_refs, d_ref_off, n, d_gcost, (gasma_cigar_op*)ctx->bm_gops.p, cap, goff))) return rc;
    if ((rc = reserve(ctx, ctx->bm_noff, ob))) return rc;
    if ((rc = reserve(ctx, ctx->bm_nops, (size_t)cap * 2))) return rc;
    uint32_t* noff = (uint32_t*)ctx->bm_noff.p;
    if ((rc = nw_dev_impl(ctx, d_reads, d_read_off, d_refs, d_ref_off, n, d_gcost, d_nwpen, (gasma_cigar_op*)ctx->bm_nops.p, cap, noff))) return rc;
    return gasma_metrics_dev(ctx, d_reads, d_read_off, n, d_gcost, (const gasma_cigar_op*)ctx->bm_gops.p, goff, d_nwpen,
                             (const gasma_cigar_op*)ctx->bm_nops.p, noff, d_answers, d_counters);
}

}  // namespace

extern "C" {

int gasma_benchmark_batch_dev(gasma_ctx* ctx, const char* d_reads, const uint32_t* d_read_off, const char* d_refs,
                              const uint32_t* d_ref_off, int64_t n_pairs, const int32_t* d_answers, uint64_t* d_counters,
                              int32_t* d_greedy_cost, int32_t* d_nw_penalty) {
    return benchmark_dev_impl(ctx, d_reads, d_read_off, d_refs, d_ref_off, n_pairs, d_answers, d_counters, d_greedy_cost,
                              d_nw_penalty);
}

int gasma_benchmark_batch(gasma_ctx* ctx, const char* reads, const uint32_t* read_off, const char* refs,
                          const uint32_t* ref_off, int64_t n, const int32_t* answers, uint64_t counters[4],
                          int32_t* greedy_cost, int32_t* nw_penalty, double device_ms[2]) {
    if (!ctx) return GASMA_E_ARG;
    if (n < 0 || !read_off || !ref_off || !counters) return fail(ctx, GASMA_E_ARG, "bad argument");
    if (n == 0) return GASMA_OK;
    int rc;
    if ((rc = check_offsets(ctx, read_off, n, "read"))) return rc;
    if ((rc = check_offsets(ctx, ref_off, n, "ref"))) return rc;
    CU(cudaSetDevice(ctx->device));
    const size_t rb = read_off[n], fb = ref_off[n], ob = (size_t)(n + 1) * 4;
    if ((rc = reserve(ctx, ctx->d_reads, rb + 16))) return rc;
    if ((rc = reserve(ctx, ctx->d_refs, fb + 16))) return rc;
    if ((rc = reserve(ctx, ctx->d_read_off, ob))) return rc;
    if ((rc = reserve(ctx, ctx->d_ref_off, ob))) return rc;
    if ((rc = reserve(ctx, ctx->bm_gcost, (size_t)n * 4))) return rc;
    if ((rc = reserve(ctx, ctx->bm_nwpen, (size_t)n * 4))) return rc;
    if ((rc = reserve(ctx, ctx->bm_counters, 32))) return rc;
    if (answers && (rc = reserve(ctx, ctx->bm_answers, (size_t)n * 4))) return rc;
    if (rb) CU(cudaMemcpyAsync(ctx->d_reads.p, reads, rb, cudaMemcpyHostToDevice, ctx->stream));
    if (fb) CU(cudaMemcpyAsync(ctx->d_refs.p, refs, fb, cudaMemcpyHostToDevice, ctx->stream));
    CU(cudaMemcpyAsync(ctx->d_read_off.p, read_off, ob, cudaMemcpyHostToDevice, ctx->stream));
    CU(cudaMemcpyAsync(ctx->d_ref_off.p, ref_off, ob, cudaMemcpyHostToDevice, ctx->stream));
    if (answers) CU(cudaMemcpyAsync(ctx->bm_answers.p, answers, (size_t)n * 4, cudaMemcpyHostToDevice, ctx->stream));
    CU(cudaMemsetAsync(ctx->bm_counters.p, 0, 32, ctx->stream));
    rc = benchmark_dev_impl(ctx, (const char*)ctx->d_reads.p, (const uint32_t*)ctx->d_read_off.p, (const char*)ctx->d_refs.p,
                            (const uint32_t*)ctx->d_ref_off.p, n, answers ? (const int32_t*)ctx->bm_answers.p : nullptr,
                            (uint64_t*)ctx->bm_counters.p, (int32_t*)ctx->bm_gcost.p, (int32_t*)ctx->bm_nwpen.p);
    if (rc) return rc;
    uint64_t* hc = reinterpret_cast<uint64_t*>(ctx->h_misc + 10);   // 8-byte aligned inside the pinned block
    CU(cudaMemcpyAsync(hc, ctx->bm_counters.p, 32, cudaMemcpyDeviceToHost, ctx->stream));
    if (greedy_cost) CU(cudaMemcpyAsync(greedy_cost, ctx->bm_gcost.p, (size_t)n * 4, cudaMemcpyDeviceToHost, ctx->stream));
    if (nw_penalty) CU(cudaMemcpyAsync(nw_penalty, ctx->bm_nwpen.p, (size_t)n * 4, cudaMemcpyDeviceToHost, ctx->stream));
    if ((rc = gasma_synchronize(ctx))) return rc;
    for (int i = 0; i < 4; ++i) counters[i] += hc[i];
    if (device_ms) {
        device_ms[0] += last_kernel_ms(ctx, 0);
        device_ms[1] += last_kernel_ms(ctx, 1);
    }
    return GASMA_OK;
}

int gasma_generate(gasma_ctx* ctx, const gasma_gen_config* cfg, uint64_t first_pair, int64_t n, char* reads,
                   uint32_t* read_off, char* refs, uint32_t* ref_off) {
    if (!ctx) return GASMA_E_ARG;
    if (!cfg || n < 0 || !reads || !read_off || !refs || !ref_off) return fail(ctx, GASMA_E_ARG, "bad argument");
    if (n == 0) { read_off[0] = ref_off[0] = 0; return GASMA_OK; }
    const int64_t fcap = gasma_generate_ref_capacity(cfg, n);
    if (fcap < 0) return fail(ctx, GASMA_E_ARG, "bad generator config");
    CU(cudaSetDevice(ctx->device));
    int rc;
    const size_t rb = (size_t)n * (size_t)cfg->length, ob = (size_t)(n + 1) * 4;
    if ((rc = reserve(ctx, ctx->d_reads, rb + 16))) return rc;
    if ((rc = reserve(ctx, ctx->d_refs, (size_t)fcap + 16))) return rc;
    if ((rc = reserve(ctx, ctx->d_read_off, ob))) return rc;
    if ((rc = reserve(ctx, ctx->d_ref_off, ob))) return rc;
    rc = gasma_generate_dev(ctx, cfg, first_pair, n, (char*)ctx->d_reads.p, (uint32_t*)ctx->d_read_off.p, (char*)ctx->d_refs.p,
                            (uint32_t*)ctx->d_ref_off.p);
    if (rc) return rc;
    CU(cudaMemcpyAsync(read_off, ctx->d_read_off.p, ob, cudaMemcpyDeviceToHost, ctx->stream));
    CU(cudaMemcpyAsync(ref_off, ctx->d_ref_off.p, ob, cudaMemcpyDeviceToHost, ctx->stream));
    if ((rc = gasma_synchronize(ctx))) return rc;
    if (rb) CU(cudaMemcpyAsync(reads, ctx->d_reads.p, rb, cudaMemcpyDeviceToHost, ctx->stream));
    if (ref_off[n]) CU(cudaMemcpyAsync(refs, ctx->d_refs.p, ref_off[n], cudaMemcpyDeviceToHost, ctx->stream));
    CU(cudaStreamSynchronize(ctx->stream));
    return GASMA_OK;
}

}  // extern "C"
