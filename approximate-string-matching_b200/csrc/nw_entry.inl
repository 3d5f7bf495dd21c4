// nw_entry.inl -- placeholder until nw.cuh / metrics.cuh land (same commit series).
extern "C" {
int gasma_nw_batch(gasma_ctx* ctx, const char*, const uint32_t*, const char*, const uint32_t*, int64_t, const int32_t*,
                   int32_t*, gasma_cigar_op*, int64_t, uint32_t*) {
    return fail(ctx, GASMA_E_UNSUPPORTED, "gasma_nw_batch: not built yet");
}
int gasma_nw_batch_dev(gasma_ctx* ctx, const char*, const uint32_t*, const char*, const uint32_t*, int64_t,
                       const int32_t*, int32_t*, gasma_cigar_op*, int64_t, uint32_t*) {
    return fail(ctx, GASMA_E_UNSUPPORTED, "gasma_nw_batch_dev: not built yet");
}
int gasma_metrics_dev(gasma_ctx* ctx, const char*, const uint32_t*, int64_t, const int32_t*, const gasma_cigar_op*,
                      const uint32_t*, const int32_t*, const gasma_cigar_op*, const uint32_t*, const int32_t*, uint64_t*) {
    return fail(ctx, GASMA_E_UNSUPPORTED, "gasma_metrics_dev: not built yet");
}
}
