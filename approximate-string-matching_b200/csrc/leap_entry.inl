// leap_entry.inl -- kernel and C-ABI entry points of the LEAP comparator; included by gasma_b200.cu.
namespace gasma {

template <int W>
__host__ __device__ constexpr size_t leap_stream_bytes(int tile) {
    return 4 * (size_t)stream_elems<W>(tile) * 2 + 2 * (size_t)(tile + 1) * 4 + 64;
}

// One pair per thread; lane masks and the level rings live in shared memory, word i of a pair at [i][thread].
template <int W, int T>
__global__ void __launch_bounds__(T)
leap_kernel(int k, int x, int o, int e, int depth, const uint8_t* __restrict__ reads, const uint32_t* __restrict__ read_off,
            const uint8_t* __restrict__ refs, const uint32_t* __restrict__ ref_off, long long n, int32_t* __restrict__ penalty,
            uint32_t* __restrict__ status) {
    constexpr int BITS = 64 * W, NW = 2 * W;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    uint32_t* s_words = reinterpret_cast<uint32_t*>(smem_raw);
    const int per_pair = leap_words_per_pair(k, NW, depth);
    uint16_t* s_r0 = reinterpret_cast<uint16_t*>(s_words + (size_t)per_pair * T);
    uint16_t* s_r1 = s_r0 + stream_elems<W>(T);
    uint16_t* s_f0 = s_r1 + stream_elems<W>(T);
    uint16_t* s_f1 = s_f0 + stream_elems<W>(T);
    uint32_t* s_roff = reinterpret_cast<uint32_t*>(s_f1 + stream_elems<W>(T));
    uint32_t* s_foff = s_roff + (T + 1);

    const int tid = threadIdx.x;
    const long long p0 = (long long)blockIdx.x * T;
    const int np = (int)((n - p0) < (long long)T ? (n - p0) : (long long)T);
    for (int i = tid; i <= np; i += T) { s_roff[i] = read_off[p0 + i]; s_foff[i] = ref_off[p0 + i]; }
    for (int i = tid; i < stream_elems<W>(T); i += T) { s_r0[i] = 0; s_r1[i] = 0; s_f0[i] = 0; s_f1[i] = 0; }
    __syncthreads();
    const uint32_t rbase = s_roff[0] & ~15u, fbase = s_foff[0] & ~15u;
    bool ok = stage_tile(reads, rbase, s_roff[np], read_off[n], s_r0, s_r1, stream_chunks<W>(T), tid, T);
    ok &= stage_tile(refs, fbase, s_foff[np], ref_off[n], s_f0, s_f1, stream_chunks<W>(T), tid, T);
    __syncthreads();

    uint32_t st = 0;
    if (tid < np) {
        const long long p = p0 + tid;
        const int m = (int)(s_roff[tid + 1] - s_roff[tid]);
        const int nn = (int)(s_foff[tid + 1] - s_foff[tid]);
        if (!ok) st |= kStatusTile;
        if (m > BITS || nn > BITS) st |= kStatusLength;
        int pen = INT32_MIN;
        if (st == 0) {
            uint64_t A0[W], A1[W], B0[W], B1[W];
            const int rbit = (int)(s_roff[tid] - rbase), fbit = (int)(s_foff[tid] - fbase);
            extract_plane<W>(reinterpret_cast<const uint32_t*>(s_r0), rbit, m, A0);
            extract_plane<W>(reinterpret_cast<const uint32_t*>(s_r1), rbit, m, A1);
            extract_plane<W>(reinterpret_cast<const uint32_t*>(s_f0), fbit, nn, B0);
            extract_plane<W>(reinterpret_cast<const uint32_t*>(s_f1), fbit, nn, B1);
            const int nl = (2 * k + 1) * NW, nr = (2 * k + 3) * depth;
            LeapStore S{s_words + tid, reinterpret_cast<int*>(s_words + (size_t)nl * T + tid),
                        reinterpret_cast<int*>(s_words + (size_t)(nl + nr) * T + tid),
                        reinterpret_cast<int*>(s_words + (size_t)(nl + 2 * nr) * T + tid), T, depth};
            pen = leap_pair<W>(k, x, o, e, A0, A1, B0, B1, m, nn, S);
        }
        penalty[p] = pen;
    }
    const uint32_t wst = __reduce_or_sync(0xffffffffu, st);
    if (wst && (tid & 31) == 0) atomicOr(status, wst);
}

}  // namespace gasma

namespace {

template <int W, int T>
int launch_leap_wt(gasma_ctx* ctx, const uint8_t* reads, const uint32_t* roff, const uint8_t* refs, const uint32_t* foff, int64_t n,
                   int32_t* penalty, int depth, size_t smem) {
    auto kern = leap_kernel<W, T>;
    CU(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    kern<<<(unsigned)((n + T - 1) / T), T, smem, ctx->stream>>>(ctx->kp.k, ctx->kp.x, ctx->kp.o, ctx->kp.e, depth, reads, roff, refs,
                                                               foff, (long long)n, penalty, (uint32_t*)ctx->misc.p);
    CU(cudaGetLastError());
    ctx->launches++;
    return GASMA_OK;
}

template <int W>
int launch_leap_w(gasma_ctx* ctx, const uint8_t* reads, const uint32_t* roff, const uint8_t* refs, const uint32_t* foff, int64_t n,
                  int32_t* penalty) {
    const int depth = leap_depth(ctx->kp.x, ctx->kp.o, ctx->kp.e);
    const size_t per_thread = (size_t)leap_words_per_pair(ctx->kp.k, 2 * W, depth) * 4;
    const size_t budget = 200 * 1024;
    // the widest tile whose lane masks + level rings fit (band k and penalties decide the per-pair footprint)
    if (per_thread * 128 + leap_stream_bytes<W>(128) <= budget / 4)
        return launch_leap_wt<W, 128>(ctx, reads, roff, refs, foff, n, penalty, depth, per_thread * 128 + leap_stream_bytes<W>(128));
    if (per_thread * 64 + leap_stream_bytes<W>(64) <= budget / 2)
        return launch_leap_wt<W, 64>(ctx, reads, roff, refs, foff, n, penalty, depth, per_thread * 64 + leap_stream_bytes<W>(64));
    if (per_thread * 32 + leap_stream_bytes<W>(32) <= budget)
        return launch_leap_wt<W, 32>(ctx, reads, roff, refs, foff, n, penalty, depth, per_thread * 32 + leap_stream_bytes<W>(32));
    return fail(ctx, GASMA_E_UNSUPPORTED, "LEAP comparator: band k=%d with these penalties does not fit shared memory", ctx->kp.k);
}

int leap_dev_impl(gasma_ctx* ctx, const char* d_reads, const uint32_t* d_read_off, const char* d_refs,
                  const uint32_t* d_ref_off, int64_t n, int32_t* d_penalty) {
    if (!ctx) return GASMA_E_ARG;
    if (n < 0 || n > (int64_t)1 << 31) return fail(ctx, GASMA_E_ARG, "n_pairs out of range");
    if (!d_read_off || !d_ref_off || !d_penalty) return fail(ctx, GASMA_E_ARG, "null offsets/penalty pointer");
    if (!aligned16(d_reads) || !aligned16(d_refs)) return fail(ctx, GASMA_E_ARG, "sequence buffers must be 16-byte aligned");
    if (ctx->W > 4) return fail(ctx, GASMA_E_UNSUPPORTED, "the LEAP comparator covers max_length <= 256 (its own _MAX_LENGTH_)");
    if (ctx->kp.o < ctx->kp.e || ctx->kp.e < 1 || ctx->kp.x < 1 || imax(ctx->kp.x, ctx->kp.o) >= kLeapMaxDepth)
        return fail(ctx, GASMA_E_UNSUPPORTED, "the LEAP comparator needs 1 <= e <= o and x, o <= %d", kLeapMaxDepth - 1);
    CU(cudaSetDevice(ctx->device));
    if (n == 0) return GASMA_OK;
    const uint8_t* r = (const uint8_t*)d_reads;
    const uint8_t* f = (const uint8_t*)d_refs;
    cudaEvent_t* evp = ctx->ev[2][ctx->ev_count[2] % gasma_ctx::kRing];
    CU(cudaEventRecord(evp[0], ctx->stream));
    int rc;
    switch (ctx->W) {
        case 1: rc = launch_leap_w<1>(ctx, r, d_read_off, f, d_ref_off, n, d_penalty); break;
        case 2: rc = launch_leap_w<2>(ctx, r, d_read_off, f, d_ref_off, n, d_penalty); break;
        case 3: rc = launch_leap_w<3>(ctx, r, d_read_off, f, d_ref_off, n, d_penalty); break;
        default: rc = launch_leap_w<4>(ctx, r, d_read_off, f, d_ref_off, n, d_penalty); break;
    }
    if (rc) return rc;
    CU(cudaEventRecord(evp[1], ctx->stream));
    ctx->ev_count[2]++;
    return GASMA_OK;
}

}  // namespace

extern "C" {

int gasma_leap_batch_dev(gasma_ctx* ctx, const char* d_reads, const uint32_t* d_read_off, const char* d_refs,
                         const uint32_t* d_ref_off, int64_t n_pairs, int32_t* d_penalty) {
    return leap_dev_impl(ctx, d_reads, d_read_off, d_refs, d_ref_off, n_pairs, d_penalty);
}

int gasma_leap_batch(gasma_ctx* ctx, const char* reads, const uint32_t* read_off, const char* refs, const uint32_t* ref_off,
                     int64_t n, int32_t* penalty) {
    if (!ctx) return GASMA_E_ARG;
    if (n < 0 || !read_off || !ref_off || !penalty) return fail(ctx, GASMA_E_ARG, "bad argument");
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
    if ((rc = reserve(ctx, ctx->d_cost, (size_t)n * 4))) return rc;
    if (rb) CU(cudaMemcpyAsync(ctx->d_reads.p, reads, rb, cudaMemcpyHostToDevice, ctx->stream));
    if (fb) CU(cudaMemcpyAsync(ctx->d_refs.p, refs, fb, cudaMemcpyHostToDevice, ctx->stream));
    CU(cudaMemcpyAsync(ctx->d_read_off.p, read_off, ob, cudaMemcpyHostToDevice, ctx->stream));
    CU(cudaMemcpyAsync(ctx->d_ref_off.p, ref_off, ob, cudaMemcpyHostToDevice, ctx->stream));
    rc = leap_dev_impl(ctx, (const char*)ctx->d_reads.p, (const uint32_t*)ctx->d_read_off.p, (const char*)ctx->d_refs.p,
                       (const uint32_t*)ctx->d_ref_off.p, n, (int32_t*)ctx->d_cost.p);
    if (rc) return rc;
    CU(cudaMemcpyAsync(penalty, ctx->d_cost.p, (size_t)n * 4, cudaMemcpyDeviceToHost, ctx->stream));
    return gasma_synchronize(ctx);
}

}  // extern "C"
