// noalloc_entry.inl -- gasma_greedy_batch_dev_noalloc: the greedy step (K1+K2+K3 + CIGAR CSR) for callers that
// cannot call cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it).  Included by
// gasma_b200.cu.  No counterpart in the reference.
//
// gasma_greedy_batch_dev keeps its scratch (per-pair CIGAR staging rows, op counts, per-tile sums / bases, the chunk
// counter) in context buffers that grow with cudaMalloc.  This variant keeps the same scratch in ONE file-scope
// __device__ array of fixed size: every GPU that loads the library holds it (8.8 MB) until the process exits, and
// the call itself allocates nothing -- no cudaMalloc / cudaFree / reserve().
//
// In host code `g_noalloc_scratch` is only a host-side stand-in for the symbol, NOT a device address: the address
// the kernels get comes from cudaGetSymbolAddress, looked up once per device with that device current.
//
// Every call on a device shares the array and it keeps the previous call's contents: the words that must start at
// zero (per-tile sums, chunk counter) are cleared on the context's stream at the start of every call.  Two calls
// that overlap on the GPU (different streams / contexts of one device) race on it -- callers serialise them.
namespace {

constexpr int kNoallocMaxPairs = GASMA_NOALLOC_MAX_PAIRS;
constexpr int kNoallocMaxLength = GASMA_NOALLOC_MAX_LENGTH;
constexpr int kNoallocTiles = (kNoallocMaxPairs + kCsrTile - 1) / kCsrTile;
constexpr size_t kNoallocStageBytes = (size_t)kNoallocMaxPairs * (2 * kNoallocMaxLength + 8) * sizeof(uint16_t);   // overflow rows
constexpr size_t kNoallocRowsOff = kNoallocStageBytes;                                                    // primary rows [pairs][kRowOps]
constexpr size_t kNoallocNopsOff = kNoallocRowsOff + (size_t)kNoallocMaxPairs * kRowPitch * sizeof(uint16_t);   // uint32 [pairs]
constexpr size_t kNoallocSumsOff = kNoallocNopsOff + (size_t)kNoallocMaxPairs * 4;      // uint32 [tiles]
constexpr size_t kNoallocWordsOff = kNoallocSumsOff + (size_t)kNoallocTiles * 4;        // [0] chunk counter, [1] op total
constexpr size_t kNoallocBytes = (kNoallocWordsOff + 64 + 255) / 256 * 256;

}  // namespace

static __device__ __align__(256) unsigned char g_noalloc_scratch[kNoallocBytes];

namespace {

constexpr int kNoallocMaxDevices = 64;
unsigned char* g_noalloc_ptr[kNoallocMaxDevices] = {};   // device address of the array, per device (0 = not looked up yet)

// once per device, with that device current
int noalloc_scratch(gasma_ctx* ctx, unsigned char** out) {
    if (ctx->device < 0 || ctx->device >= kNoallocMaxDevices) return fail(ctx, GASMA_E_ARG, "device index %d out of range", ctx->device);
    if (!g_noalloc_ptr[ctx->device]) {
        void* p = nullptr;
        CU(cudaSetDevice(ctx->device));
        CU(cudaGetSymbolAddress(&p, g_noalloc_scratch));
        if (!p) return fail(ctx, GASMA_E_CUDA, "cudaGetSymbolAddress returned a null address for the static scratch");
        g_noalloc_ptr[ctx->device] = (unsigned char*)p;
    }
    *out = g_noalloc_ptr[ctx->device];
    return GASMA_OK;
}

}  // namespace

extern "C" {

int gasma_greedy_batch_dev_noalloc(gasma_ctx* ctx, const char* d_reads, const uint32_t* d_read_off, const char* d_refs,
                                   const uint32_t* d_ref_off, int64_t n, int32_t* d_cost, gasma_cigar_op* d_ops,
                                   int64_t ops_cap, uint32_t* d_op_off) {
    if (!ctx) return GASMA_E_ARG;
    if (n < 0) return fail(ctx, GASMA_E_ARG, "n_pairs out of range");
    if (n > kNoallocMaxPairs)
        return fail(ctx, GASMA_E_UNSUPPORTED, "gasma_greedy_batch_dev_noalloc: %lld pairs exceed the static scratch's maximum of %d "
                    "(GASMA_NOALLOC_MAX_PAIRS); split the batch or use gasma_greedy_batch_dev", (long long)n, kNoallocMaxPairs);
    if (ctx->params.max_length > kNoallocMaxLength)
        return fail(ctx, GASMA_E_UNSUPPORTED, "gasma_greedy_batch_dev_noalloc: max_length %d exceeds the static scratch's maximum of %d "
                    "(GASMA_NOALLOC_MAX_LENGTH)", ctx->params.max_length, kNoallocMaxLength);
    if (!d_read_off || !d_ref_off || !d_cost) return fail(ctx, GASMA_E_ARG, "null offsets/cost pointer");
    if ((d_ops != nullptr) != (d_op_off != nullptr)) return fail(ctx, GASMA_E_ARG, "ops and op_off must be given together");
    if (!aligned16(d_reads) || !aligned16(d_refs)) return fail(ctx, GASMA_E_ARG, "sequence buffers must be 16-byte aligned");
    CU(cudaSetDevice(ctx->device));
    if (n == 0) {
        if (d_op_off) CU(cudaMemsetAsync(d_op_off, 0, 4, ctx->stream));
        return GASMA_OK;
    }
    unsigned char* base = nullptr;
    int rc = noalloc_scratch(ctx, &base);
    if (rc) return rc;
    const int64_t ntiles = (n + kCsrTile - 1) / kCsrTile;
    const int stride = 2 * ctx->params.max_length + 8;
    const bool want_ops = d_ops != nullptr;
    uint16_t* stage = want_ops ? (uint16_t*)base : nullptr;
    uint16_t* rows = (uint16_t*)(base + kNoallocRowsOff);   // always: the kernel stages ops unconditionally
    uint32_t* nops = want_ops ? (uint32_t*)(base + kNoallocNopsOff) : nullptr;
    uint32_t* tsums = want_ops ? (uint32_t*)(base + kNoallocSumsOff) : nullptr;
    uint32_t* words = (uint32_t*)(base + kNoallocWordsOff);
    uint32_t* misc = (uint32_t*)ctx->misc.p;   // status + step counter stay with the context (allocated by gasma_create)
    // the array keeps the last call's contents: clear what must start at zero (tile sums, chunk counter, total)
    CU(cudaMemsetAsync(base + kNoallocSumsOff, 0, kNoallocBytes - kNoallocSumsOff, ctx->stream));
    const uint8_t* r = (const uint8_t*)d_reads;
    const uint8_t* f = (const uint8_t*)d_refs;
    unsigned long long* steps = (unsigned long long*)(misc + 2);
    const int warps = greedy_tile_threads(ctx) / 32;
    // the launch helpers clear the chunk counter they are given unless the context says ITS counter is clean; ours was
    // cleared above, and the context's own counter is not touched by this call
    const bool keep_clean = ctx->counter_clean;
    ctx->counter_clean = true;
    switch (ctx->W) {
        case 1: rc = launch_greedy_w<1>(ctx, warps, r, d_read_off, f, d_ref_off, n, d_cost, rows, stage, stride, nops, words, misc, steps, tsums); break;
        case 2: rc = launch_greedy_w<2>(ctx, warps, r, d_read_off, f, d_ref_off, n, d_cost, rows, stage, stride, nops, words, misc, steps, tsums); break;
        case 3: rc = launch_greedy_w<3>(ctx, warps, r, d_read_off, f, d_ref_off, n, d_cost, rows, stage, stride, nops, words, misc, steps, tsums); break;
        default: rc = launch_greedy_w<4>(ctx, warps, r, d_read_off, f, d_ref_off, n, d_cost, rows, stage, stride, nops, words, misc, steps, tsums); break;
    }
    ctx->counter_clean = keep_clean;
    if (rc) return rc;
    if (want_ops) {
        compact_rows_kernel<kCsrTile><<<(unsigned)ntiles, kCsrTile, 0, ctx->stream>>>(
            rows, stage, stride, nops, tsums, tsums, 0, nullptr, words + 1, (long long)n, d_ops, (long long)ops_cap, d_op_off, nullptr);
        CU(cudaGetLastError());
        ctx->launches += 1;
    }
    return GASMA_OK;
}

}  // extern "C"
