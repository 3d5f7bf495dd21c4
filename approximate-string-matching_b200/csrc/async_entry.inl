// async_entry.inl -- gasma_greedy_submit / gasma_greedy_submit_packed / gasma_greedy_wait: the host-pointer greedy call
// as an asynchronous pair, two calls in flight per context.  Included by gasma_b200.cu.
//
// A synchronous host-pointer call ends with a tail in which the link is idle: the last slice's kernels, its penalties
// and CSR offsets coming back, one host round trip to learn how many CIGAR ops to fetch, then that fetch
// (GASMA_TRACE: ~0.35 ms of a 6 ms call on ASCII input, ~0.35 of 2.1 ms on packed input).  With two calls in flight
// the tail of call i runs under the input copies of call i+1.
//
// Each of the two in-flight slots is a full set of device mirrors, streams and events -- a child context created on
// first use with the parent's parameters and settings -- and the call itself is the ordinary synchronous entry point
// run on the child by a host thread of its own (the call needs the host between its stages: offsets are validated
// behind the first DMA, the op fetch waits for the offsets).  Nothing about the results changes: a submitted call IS
// gasma_greedy_batch / gasma_greedy_batch_packed on identical arguments.
//
// No counterpart in the reference (its loop is synchronous, benchmark_utils.h:238-252); this is how a caller that
// streams batches through the library keeps PCIe busy.
#include <future>

struct gasma_async_slot {
    gasma_ctx* child = nullptr;
    std::future<int> result;
    bool busy = false;
    int ticket = -1;
    int64_t launches_before = 0;
};

struct gasma_async_state {
    gasma_async_slot slot[2];
    int next_ticket = 0;
};

namespace {

// the child runs with the parent's CURRENT settings (band, widened band, mask variant), checked at every submit
int async_child(gasma_ctx* ctx, int s, gasma_ctx** out) {
    if (!ctx->async) ctx->async = new gasma_async_state();
    gasma_async_slot& sl = ctx->async->slot[s];
    if (!sl.child) {
        gasma_ctx* c = nullptr;
        gasma_params p = ctx->params;
        const int rc = gasma_create(&p, ctx->device, &c);
        if (rc) return fail(ctx, rc, "creating the mirror context of an asynchronous call failed: %s", gasma_last_error(nullptr));
        sl.child = c;
    }
    gasma_ctx* c = sl.child;
    if (c->params.k != ctx->params.k || c->correction != ctx->correction || c->max_delta != ctx->max_delta) {
        int rc = gasma_set_correction(c, 0, 0);
        if (!rc) rc = gasma_set_mask_variant(c, 0);
        if (!rc) rc = gasma_set_band(c, ctx->params.k);
        if (!rc && ctx->correction) rc = gasma_set_correction(c, 1, ctx->max_delta);
        if (rc) return fail(ctx, rc, "%s", c->err.c_str());
    }
    if (c->mask_variant != ctx->mask_variant) {
        const int rc = gasma_set_mask_variant(c, ctx->mask_variant);
        if (rc) return fail(ctx, rc, "%s", c->err.c_str());
    }
    *out = c;
    return GASMA_OK;
}

int async_pick_slot(gasma_ctx* ctx, int* ticket, int* slot) {
    if (!ticket) return fail(ctx, GASMA_E_ARG, "ticket is null");
    *ticket = -1;
    if (!ctx->async) ctx->async = new gasma_async_state();
    for (int s = 0; s < 2; ++s)
        if (!ctx->async->slot[s].busy) { *slot = s; return GASMA_OK; }
    return fail(ctx, GASMA_E_UNSUPPORTED, "two asynchronous calls are in flight (tickets %d and %d): gasma_greedy_wait one of them first",
                ctx->async->slot[0].ticket, ctx->async->slot[1].ticket);
}

template <class F>
int async_start(gasma_ctx* ctx, int s, int* ticket, F&& call) {
    gasma_async_slot& sl = ctx->async->slot[s];
    sl.launches_before = sl.child->launches;
    try {
        sl.result = std::async(std::launch::async, std::forward<F>(call));
    } catch (const std::exception& ex) {
        return fail(ctx, GASMA_E_INTERNAL, "could not start the host thread of an asynchronous call: %s", ex.what());
    }
    sl.busy = true;
    sl.ticket = ctx->async->next_ticket++;
    *ticket = sl.ticket;
    return GASMA_OK;
}

void async_destroy(gasma_ctx* ctx) {
    if (!ctx->async) return;
    for (auto& sl : ctx->async->slot) {
        if (sl.busy && sl.result.valid()) sl.result.wait();   // the caller's buffers are in use until the call has ended
        if (sl.child) gasma_destroy(sl.child);
    }
    delete ctx->async;
    ctx->async = nullptr;
}

}  // namespace

extern "C" {

int gasma_greedy_submit(gasma_ctx* ctx, const char* reads, const uint32_t* read_off, const char* refs, const uint32_t* ref_off,
                        int64_t n, int32_t* cost, gasma_cigar_op* ops, int64_t ops_cap, uint32_t* op_off, int* ticket) {
    if (!ctx) return GASMA_E_ARG;
    int s = 0, rc = async_pick_slot(ctx, ticket, &s);
    if (rc) return rc;
    gasma_ctx* c = nullptr;
    if ((rc = async_child(ctx, s, &c))) return rc;
    return async_start(ctx, s, ticket, [=] { return gasma_greedy_batch(c, reads, read_off, refs, ref_off, n, cost, ops, ops_cap, op_off); });
}

int gasma_greedy_submit_packed(gasma_ctx* ctx, const uint16_t* read_p0, const uint16_t* read_p1, const uint32_t* read_off,
                               const uint16_t* ref_p0, const uint16_t* ref_p1, const uint32_t* ref_off, int64_t n, int32_t* cost,
                               gasma_cigar_op* ops, int64_t ops_cap, uint32_t* op_off, int* ticket) {
    if (!ctx) return GASMA_E_ARG;
    int s = 0, rc = async_pick_slot(ctx, ticket, &s);
    if (rc) return rc;
    gasma_ctx* c = nullptr;
    if ((rc = async_child(ctx, s, &c))) return rc;
    return async_start(ctx, s, ticket, [=] {
        return gasma_greedy_batch_packed(c, read_p0, read_p1, read_off, ref_p0, ref_p1, ref_off, n, cost, ops, ops_cap, op_off);
    });
}

int gasma_greedy_wait(gasma_ctx* ctx, int ticket) {
    if (!ctx) return GASMA_E_ARG;
    if (ctx->async)
        for (auto& sl : ctx->async->slot) {
            if (!sl.busy || sl.ticket != ticket) continue;
            int rc = GASMA_E_INTERNAL;
            try {
                rc = sl.result.get();
            } catch (const std::exception& ex) {
                sl.child->err = ex.what();
            }
            sl.busy = false;
            ctx->launches += sl.child->launches - sl.launches_before;
            if (rc) return fail(ctx, rc, "%s", sl.child->err.c_str());
            return GASMA_OK;
        }
    return fail(ctx, GASMA_E_ARG, "ticket %d is not in flight on this context", ticket);
}

}  // extern "C"
