// nccl_entry.inl -- gasma_allreduce_counters: the path's ONLY collective, callable from C++ without torch.
//
// benchmark::_run_benchmark keeps four counters (total_tests, nw_correct, greedy_correct, greedy_coverage;
// benchmark_utils.h:238,253-258).  When the pairs of a run are sharded over the GPUs of a box every context ends with
// its own counts in device memory; their sum over NVLink 5 / NVSwitch is one ncclAllReduce of `count` uint64 words
// (32 bytes for the four counters: latency only).  Included by gasma_b200.cu.
//
// NCCL is bound at RUN time (dlopen of libnccl.so.2; the image has 2.27.3 system-wide and torch bundles its own):
// libgasma_b200.so has no link-time NCCL dependency, a process that already holds a copy (python with torch) gets that
// copy, and single-GPU users never load it.  One communicator clique per device list, cached for the process.
#include <dlfcn.h>

#include <mutex>
#include <vector>

namespace {

struct NcclApi {
    void* handle = nullptr;
    // the handful of entry points used, with nccl.h's signatures (ncclComm_t is an opaque pointer, results are ints)
    int (*CommInitAll)(void** comms, int ndev, const int* devlist) = nullptr;
    int (*CommDestroy)(void* comm) = nullptr;
    int (*GroupStart)() = nullptr;
    int (*GroupEnd)() = nullptr;
    int (*AllReduce)(const void* send, void* recv, size_t count, int datatype, int op, void* comm, cudaStream_t stream) = nullptr;
    const char* (*GetErrorString)(int) = nullptr;
    int (*GetVersion)(int*) = nullptr;
    std::string why;
};
constexpr int kNcclUint64 = 5, kNcclSum = 0;   // nccl.h: ncclUint64, ncclSum

NcclApi& nccl_api() {
    static NcclApi api;
    static std::once_flag once;
    std::call_once(once, [] {
        const char* names[] = {getenv("GASMA_NCCL_LIB"), "libnccl.so.2", "libnccl.so"};
        for (const char* n : names) {
            if (!n) continue;
            api.handle = dlopen(n, RTLD_NOW | RTLD_GLOBAL);
            if (api.handle) break;
        }
        if (!api.handle) { api.why = std::string("libnccl.so.2 not found (") + (dlerror() ? dlerror() : "?") + ")"; return; }
        auto sym = [&](const char* s) { void* p = dlsym(api.handle, s); if (!p) api.why = std::string("missing symbol ") + s; return p; };
        api.CommInitAll = reinterpret_cast<decltype(api.CommInitAll)>(sym("ncclCommInitAll"));
        api.CommDestroy = reinterpret_cast<decltype(api.CommDestroy)>(sym("ncclCommDestroy"));
        api.GroupStart = reinterpret_cast<decltype(api.GroupStart)>(sym("ncclGroupStart"));
        api.GroupEnd = reinterpret_cast<decltype(api.GroupEnd)>(sym("ncclGroupEnd"));
        api.AllReduce = reinterpret_cast<decltype(api.AllReduce)>(sym("ncclAllReduce"));
        api.GetErrorString = reinterpret_cast<decltype(api.GetErrorString)>(sym("ncclGetErrorString"));
        api.GetVersion = reinterpret_cast<decltype(api.GetVersion)>(sym("ncclGetVersion"));
    });
    return api;
}

struct NcclClique {
    std::vector<int> devices;
    std::vector<void*> comms;
};
std::mutex g_nccl_mutex;
std::vector<NcclClique> g_nccl_cliques;   // kept until the process exits (communicators are expensive to build)

}  // namespace

extern "C" {

int gasma_nccl_version(void) {
    NcclApi& api = nccl_api();
    int v = 0;
    if (!api.handle || !api.why.empty() || api.GetVersion(&v) != 0) return 0;
    return v;
}

int gasma_allreduce_counters(gasma_ctx* const* ctxs, int n_ctx, uint64_t* const* d_counters, int count) {
    if (!ctxs || n_ctx < 1 || !ctxs[0]) return GASMA_E_ARG;
    gasma_ctx* ctx = ctxs[0];   // errors are reported on the first context
    if (!d_counters || count < 1 || count > 1024) return fail(ctx, GASMA_E_ARG, "bad counter arrays / count");
    std::vector<int> devs((size_t)n_ctx);
    for (int i = 0; i < n_ctx; ++i) {
        if (!ctxs[i] || !d_counters[i]) return fail(ctx, GASMA_E_ARG, "null context or counter array %d", i);
        devs[(size_t)i] = ctxs[i]->device;
        for (int j = 0; j < i; ++j)
            if (devs[(size_t)j] == devs[(size_t)i])
                return fail(ctx, GASMA_E_ARG, "contexts %d and %d share device %d: NCCL wants one rank per device", j, i, devs[(size_t)i]);
    }
    NcclApi& api = nccl_api();
    if (!api.handle || !api.why.empty()) return fail(ctx, GASMA_E_UNSUPPORTED, "NCCL unavailable: %s", api.why.c_str());
    std::lock_guard<std::mutex> lock(g_nccl_mutex);
    NcclClique* cl = nullptr;
    for (auto& c : g_nccl_cliques)
        if (c.devices == devs) cl = &c;
    if (!cl) {
        NcclClique c;
        c.devices = devs;
        c.comms.assign((size_t)n_ctx, nullptr);
        const int rc = api.CommInitAll(c.comms.data(), n_ctx, devs.data());
        if (rc != 0) return fail(ctx, GASMA_E_CUDA, "ncclCommInitAll over %d devices failed: %s", n_ctx, api.GetErrorString(rc));
        g_nccl_cliques.push_back(c);
        cl = &g_nccl_cliques.back();
    }
    // one in-place all-reduce per rank, on each context's own stream (ordered after the kernels that produced the counts)
    int rc = api.GroupStart();
    for (int i = 0; i < n_ctx && rc == 0; ++i) {
        CU(cudaSetDevice(ctxs[i]->device));
        rc = api.AllReduce(d_counters[i], d_counters[i], (size_t)count, kNcclUint64, kNcclSum, cl->comms[(size_t)i], ctxs[i]->stream);
    }
    const int rc2 = api.GroupEnd();
    if (rc != 0 || rc2 != 0) return fail(ctx, GASMA_E_CUDA, "ncclAllReduce failed: %s", api.GetErrorString(rc ? rc : rc2));
    for (int i = 0; i < n_ctx; ++i) {
        CU(cudaSetDevice(ctxs[i]->device));
        CU(cudaStreamSynchronize(ctxs[i]->stream));
    }
    CU(cudaSetDevice(ctx->device));
    return GASMA_OK;
}

// Host-array flavour for C++ callers without CUDA headers (host/benchmark_gpu.h): counters[i] = `count` uint64 words in
// HOST memory holding context i's counts; they travel to a scratch word block of their context, are summed with
// gasma_allreduce_counters and come back, so that afterwards every host array holds the total.
int gasma_allreduce_counters_host(gasma_ctx* const* ctxs, int n_ctx, uint64_t* const* counters, int count) {
    if (!ctxs || n_ctx < 1 || !ctxs[0]) return GASMA_E_ARG;
    gasma_ctx* ctx = ctxs[0];
    if (!counters || count < 1 || count > 1024) return fail(ctx, GASMA_E_ARG, "bad counter arrays / count");
    std::vector<uint64_t*> dev((size_t)n_ctx);
    for (int i = 0; i < n_ctx; ++i) {
        if (!ctxs[i] || !counters[i]) return fail(ctx, GASMA_E_ARG, "null context or counter array %d", i);
        CU(cudaSetDevice(ctxs[i]->device));
        const int rc = reserve(ctxs[i], ctxs[i]->nccl_words, (size_t)count * 8);
        if (rc) return fail(ctx, rc, "%s", ctxs[i]->err.c_str());
        dev[(size_t)i] = (uint64_t*)ctxs[i]->nccl_words.p;
        CU(cudaMemcpyAsync(dev[(size_t)i], counters[i], (size_t)count * 8, cudaMemcpyHostToDevice, ctxs[i]->stream));
    }
    const int rc = gasma_allreduce_counters(ctxs, n_ctx, dev.data(), count);
    if (rc) return rc;
    for (int i = 0; i < n_ctx; ++i) {
        CU(cudaSetDevice(ctxs[i]->device));
        CU(cudaMemcpyAsync(counters[i], dev[(size_t)i], (size_t)count * 8, cudaMemcpyDeviceToHost, ctxs[i]->stream));
        CU(cudaStreamSynchronize(ctxs[i]->stream));
    }
    CU(cudaSetDevice(ctx->device));
    return GASMA_OK;
}

}  // extern "C"
