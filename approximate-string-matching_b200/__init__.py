"""gasma_b200 -- Python plumbing over the C ABI of libgasma_b200.so (include/gasma_b200.h).

The product is the CUDA library; this module only loads it with ctypes, owns a context and passes
pointers (numpy host arrays for the host entry points, torch CUDA tensors' data_ptr() for the _dev
entry points).  There is no Python or CPU compute path: if the library or a B200 is missing, calls
raise.  The directory name contains a '-', so import it through `load()` in __graft_entry__.py or
tests/conftest.py (importlib by path) under the module name `gasma_b200`.
"""
import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("GASMA_B200_LIB") or os.path.join(HERE, "libgasma_b200.so")   # override: kernel A/B experiments

GLOBAL, SEMI_GLOBAL = 0, 1
OP_LETTERS = "MID=X"
MAX_K = 50


class GasmaError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__("gasma error %d: %s" % (code, msg))
        self.code = code


class Params(C.Structure):
    _fields_ = [("alignment_type", C.c_int32), ("x", C.c_int32), ("o", C.c_int32), ("e", C.c_int32),
                ("k", C.c_int32), ("max_length", C.c_int32),
                ("match_prob", C.c_double), ("mismatch_prob", C.c_double), ("indel_prob", C.c_double)]


class GenConfig(C.Structure):
    _fields_ = [("seed", C.c_uint64), ("length", C.c_int32), ("mode", C.c_int32), ("num_errors", C.c_int32),
                ("reserved", C.c_int32), ("thr_mismatch", C.c_uint64), ("thr_sub", C.c_uint64),
                ("thr_ins", C.c_uint64), ("thr_del", C.c_uint64)]


# every symbol include/gasma_b200.h declares (tests check that the library exports all of them)
EXPORTS = [
    "gasma_default_params", "gasma_create", "gasma_destroy", "gasma_last_error", "gasma_set_band",
    "gasma_set_correction", "gasma_set_stream", "gasma_synchronize", "gasma_launch_count", "gasma_kernel_time_ms", "gasma_last_step_count",
    "gasma_pack2bit_batch", "gasma_pack2bit_batch_dev", "gasma_greedy_batch", "gasma_greedy_batch_dev",
    "gasma_nw_batch", "gasma_nw_batch_dev", "gasma_metrics_dev",
    "gasma_generate_ref_capacity", "gasma_generate_dev", "gasma_generate", "gasma_render_cigar",
    "gasma_benchmark_batch", "gasma_benchmark_batch_dev", "gasma_leap_batch", "gasma_leap_batch_dev",
    "gasma_greedy_batch_dev_noalloc", "gasma_set_mask_variant",
    "gasma_pack2bit_stream", "gasma_pack2bit_stream_dev", "gasma_greedy_batch_packed", "gasma_greedy_batch_packed_dev",
    "gasma_allreduce_counters", "gasma_allreduce_counters_host", "gasma_nccl_version",
    "gasma_greedy_submit", "gasma_greedy_submit_packed", "gasma_greedy_wait",
]

_lib = None


def lib():
    """Load libgasma_b200.so; fail loudly if the CUDA extension has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError("libgasma_b200.so is missing (%s): build it with __graft_entry__.build() / "
                          "python approximate-string-matching_b200/build.py -- there is no CPU fallback" % LIB_PATH)
    L = C.CDLL(LIB_PATH)
    vp, i64, i32 = C.c_void_p, C.c_int64, C.c_int
    L.gasma_default_params.argtypes = [C.POINTER(Params)]
    L.gasma_default_params.restype = None
    L.gasma_create.argtypes = [C.POINTER(Params), i32, C.POINTER(vp)]
    L.gasma_destroy.argtypes = [vp]
    L.gasma_destroy.restype = None
    L.gasma_last_error.argtypes = [vp]
    L.gasma_last_error.restype = C.c_char_p
    L.gasma_set_band.argtypes = [vp, i32]
    L.gasma_set_correction.argtypes = [vp, i32, i32]
    L.gasma_set_mask_variant.argtypes = [vp, i32]
    L.gasma_set_stream.argtypes = [vp, vp]
    L.gasma_synchronize.argtypes = [vp]
    L.gasma_launch_count.argtypes = [vp]
    L.gasma_launch_count.restype = i64
    L.gasma_kernel_time_ms.argtypes = [vp, i32, C.POINTER(i64)]
    L.gasma_kernel_time_ms.restype = C.c_double
    L.gasma_last_step_count.argtypes = [vp]
    L.gasma_last_step_count.restype = i64
    for name in ("gasma_pack2bit_batch", "gasma_pack2bit_batch_dev"):
        getattr(L, name).argtypes = [vp, vp, vp, i64, vp, vp]
    for name in ("gasma_greedy_batch", "gasma_greedy_batch_dev", "gasma_greedy_batch_dev_noalloc"):
        getattr(L, name).argtypes = [vp, vp, vp, vp, vp, i64, vp, vp, i64, vp]
    for name in ("gasma_nw_batch", "gasma_nw_batch_dev"):
        getattr(L, name).argtypes = [vp, vp, vp, vp, vp, i64, vp, vp, vp, i64, vp]
    for name in ("gasma_pack2bit_stream", "gasma_pack2bit_stream_dev"):
        getattr(L, name).argtypes = [vp, vp, i64, vp, vp]
    for name in ("gasma_greedy_batch_packed", "gasma_greedy_batch_packed_dev"):
        getattr(L, name).argtypes = [vp, vp, vp, vp, vp, vp, vp, i64, vp, vp, i64, vp]
    L.gasma_greedy_submit.argtypes = [vp, vp, vp, vp, vp, i64, vp, vp, i64, vp, C.POINTER(C.c_int)]
    L.gasma_greedy_submit_packed.argtypes = [vp, vp, vp, vp, vp, vp, vp, i64, vp, vp, i64, vp, C.POINTER(C.c_int)]
    L.gasma_greedy_wait.argtypes = [vp, i32]
    L.gasma_allreduce_counters.argtypes = [vp, i32, vp, i32]
    L.gasma_allreduce_counters_host.argtypes = [vp, i32, vp, i32]
    L.gasma_nccl_version.argtypes = []
    L.gasma_metrics_dev.argtypes = [vp, vp, vp, i64] + [vp] * 8
    L.gasma_generate_ref_capacity.argtypes = [C.POINTER(GenConfig), i64]
    L.gasma_generate_ref_capacity.restype = i64
    L.gasma_generate_dev.argtypes = [vp, C.POINTER(GenConfig), C.c_uint64, i64, vp, vp, vp, vp]
    L.gasma_render_cigar.argtypes = [vp, i32, vp, i32]
    L.gasma_generate.argtypes = [vp, C.POINTER(GenConfig), C.c_uint64, i64, vp, vp, vp, vp]
    L.gasma_benchmark_batch.argtypes = [vp, vp, vp, vp, vp, i64, vp, vp, vp, vp, vp]
    L.gasma_benchmark_batch_dev.argtypes = [vp, vp, vp, vp, vp, i64, vp, vp, vp, vp]
    for name in ("gasma_leap_batch", "gasma_leap_batch_dev"):
        getattr(L, name).argtypes = [vp, vp, vp, vp, vp, i64, vp]
    _lib = L
    return L


def allreduce_counters(aligners, counter_tensors):
    """gasma_allreduce_counters: the in-place NCCL sum of per-device uint64 counter arrays (torch CUDA tensors, one per
    aligner, each on its aligner's device)."""
    n = len(aligners)
    ctxs = (C.c_void_p * n)(*[a._ctx.value for a in aligners])
    ptrs = (C.c_void_p * n)(*[t.data_ptr() for t in counter_tensors])
    rc = lib().gasma_allreduce_counters(ctxs, n, ptrs, counter_tensors[0].numel())
    if rc != 0:
        raise GasmaError(rc, lib().gasma_last_error(aligners[0]._ctx).decode())


def _np_ptr(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def _dev_ptr(t):
    return None if t is None else C.c_void_p(t.data_ptr())


def gen_config(seed, length, error_rate, mismatch_rate=0.96, mode=0, profile=None):
    """Same arithmetic as the reference's Dataset for the edit count: ceil(L * e) in float
    (benchmark_dataset.h:153-156)."""
    ne = int(np.ceil(np.float32(length) * np.float32(error_rate)))
    c = GenConfig(seed, length, mode, ne, 0, int(mismatch_rate * 4294967296.0), 0, 0, 0)
    if mode == 2:
        ps, pi, pd = profile
        c.thr_del = int(pd * 4294967296.0)
        c.thr_ins = c.thr_del + int(pi * 4294967296.0)
        c.thr_sub = c.thr_ins + int(ps * 4294967296.0)
    return c


def render_ops(ops):
    return "".join("%d%s" % (int(o) >> 3, OP_LETTERS[int(o) & 7]) for o in ops)


class Aligner:
    """One gasma_ctx: the batched counterpart of one hurdle_matrix<int_128bit> object
    (GASMA/hurdle_matrix.h:552-562; benchmark_utils.h:287)."""

    def __init__(self, k=3, x=1, o=1, e=1, max_length=128, alignment_type=GLOBAL,
                 match_prob=0.80, mismatch_prob=0.20 / 3, indel_prob=0.40 / 3, device=0):
        self._L = lib()
        self.params = Params(alignment_type, x, o, e, k, max_length, match_prob, mismatch_prob, indel_prob)
        self._ctx = C.c_void_p()
        rc = self._L.gasma_create(C.byref(self.params), device, C.byref(self._ctx))
        if rc != 0:
            raise GasmaError(rc, self._L.gasma_last_error(None).decode())
        self.W = max_length // 64
        self.device = device
        self._in_flight = {}   # ticket -> arrays of a submitted call (kept alive until greedy_wait)

    def close(self):
        ctx = getattr(self, "_ctx", None)
        if ctx is not None and ctx.value:
            self._L.gasma_destroy(ctx)
            self._ctx = None

    __del__ = close

    def _check(self, rc):
        if rc != 0:
            raise GasmaError(rc, self._L.gasma_last_error(self._ctx).decode())

    # -- context ------------------------------------------------------------------------------
    def set_band(self, k):
        self._check(self._L.gasma_set_band(self._ctx, k))
        self.params.k = k

    def set_correction(self, enabled, max_delta=0):
        """The reference's -DCORRECTION band (hurdle_matrix.h:634-641): lanes widened towards lane n-m; max_delta =
        largest |ref_len - read_len| the context keeps lane slots for."""
        self._check(self._L.gasma_set_correction(self._ctx, 1 if enabled else 0, max_delta))

    def set_mask_variant(self, variant):
        """0 = flip_short_hurdles(1) (as shipped), 1 = flip_short_hurdles(2), 2 = + flip_short_matches(1) (utils.h:200-240)."""
        self._check(self._L.gasma_set_mask_variant(self._ctx, variant))

    def set_stream(self, cuda_stream_ptr):
        self._check(self._L.gasma_set_stream(self._ctx, C.c_void_p(cuda_stream_ptr)))

    def synchronize(self):
        self._check(self._L.gasma_synchronize(self._ctx))

    @property
    def launch_count(self):
        return int(self._L.gasma_launch_count(self._ctx))

    def kernel_time_ms(self, which=0):
        """(summed ms, launches) of the dominant kernel since the last call."""
        n = C.c_int64(0)
        ms = float(self._L.gasma_kernel_time_ms(self._ctx, which, C.byref(n)))
        return ms, int(n.value)

    def last_step_count(self):
        return int(self._L.gasma_last_step_count(self._ctx))

    # -- host entry points (numpy) -------------------------------------------------------------
    def pack2bit_batch(self, seqs, off):
        n = len(off) - 1
        p0 = np.zeros((n, self.W), np.uint64)
        p1 = np.zeros((n, self.W), np.uint64)
        self._check(self._L.gasma_pack2bit_batch(self._ctx, _np_ptr(seqs), _np_ptr(off), n, _np_ptr(p0), _np_ptr(p1)))
        return p0, p1

    def greedy_batch(self, reads, read_off, refs, ref_off, want_ops=True, ops_cap=None, out=None):
        n = len(read_off) - 1
        if out is None:
            out = {"cost": np.empty(n, np.int32)}
            if want_ops:
                cap = ops_cap if ops_cap is not None else max(16, 24 * n)
                out["ops"] = np.empty(cap, np.uint16)
                out["op_off"] = np.empty(n + 1, np.uint32)
        ops = out.get("ops") if want_ops else None
        op_off = out.get("op_off") if want_ops else None
        rc = self._L.gasma_greedy_batch(self._ctx, _np_ptr(reads), _np_ptr(read_off), _np_ptr(refs), _np_ptr(ref_off),
                                        n, _np_ptr(out["cost"]), _np_ptr(ops), 0 if ops is None else len(ops),
                                        _np_ptr(op_off))
        if rc == -4 and ops_cap is None and want_ops:  # GASMA_E_CAPACITY: retry once with the exact size
            out["ops"] = np.empty(int(out["op_off"][n]), np.uint16)
            return self.greedy_batch(reads, read_off, refs, ref_off, True, len(out["ops"]), out)
        self._check(rc)
        if want_ops:
            out["n_ops"] = int(out["op_off"][n])
        return out

    # -- packed input (2-bit streams: one uint16 per 16 characters and plane, see include/gasma_b200.h) ----------------
    def pack2bit_stream(self, seqs, total=None):
        total = len(seqs) if total is None else total
        nw = (total + 15) // 16
        p0, p1 = np.zeros(nw, np.uint16), np.zeros(nw, np.uint16)
        self._check(self._L.gasma_pack2bit_stream(self._ctx, _np_ptr(seqs), total, _np_ptr(p0), _np_ptr(p1)))
        return p0, p1

    def pack2bit_stream_dev(self, d_seqs, total, d_p0, d_p1):
        self._check(self._L.gasma_pack2bit_stream_dev(self._ctx, _dev_ptr(d_seqs), total, _dev_ptr(d_p0), _dev_ptr(d_p1)))

    def greedy_batch_packed(self, r0, r1, read_off, f0, f1, ref_off, want_ops=True, ops_cap=None, out=None):
        n = len(read_off) - 1
        if out is None:
            out = {"cost": np.empty(n, np.int32)}
            if want_ops:
                out["ops"] = np.empty(ops_cap if ops_cap is not None else max(16, 24 * n), np.uint16)
                out["op_off"] = np.empty(n + 1, np.uint32)
        ops = out.get("ops") if want_ops else None
        op_off = out.get("op_off") if want_ops else None
        self._check(self._L.gasma_greedy_batch_packed(self._ctx, _np_ptr(r0), _np_ptr(r1), _np_ptr(read_off), _np_ptr(f0), _np_ptr(f1),
                                                      _np_ptr(ref_off), n, _np_ptr(out["cost"]), _np_ptr(ops),
                                                      0 if ops is None else len(ops), _np_ptr(op_off)))
        if want_ops:
            out["n_ops"] = int(out["op_off"][n])
        return out

    def greedy_batch_packed_dev(self, d_r0, d_r1, d_read_off, d_f0, d_f1, d_ref_off, n, d_cost, d_ops=None, d_op_off=None):
        self._check(self._L.gasma_greedy_batch_packed_dev(self._ctx, _dev_ptr(d_r0), _dev_ptr(d_r1), _dev_ptr(d_read_off), _dev_ptr(d_f0),
                                                          _dev_ptr(d_f1), _dev_ptr(d_ref_off), n, _dev_ptr(d_cost), _dev_ptr(d_ops),
                                                          0 if d_ops is None else d_ops.numel(), _dev_ptr(d_op_off)))

    # -- asynchronous pair: two host-pointer calls in flight per context (include/gasma_b200.h) -------------------------
    def _async_out(self, n, want_ops, ops_cap, out):
        if out is None:
            out = {"cost": np.empty(n, np.int32)}
            if want_ops:
                out["ops"] = np.empty(ops_cap if ops_cap is not None else max(16, 24 * n), np.uint16)
                out["op_off"] = np.empty(n + 1, np.uint32)
        return out, (out.get("ops") if want_ops else None), (out.get("op_off") if want_ops else None)

    def greedy_submit(self, reads, read_off, refs, ref_off, want_ops=True, ops_cap=None, out=None):
        """Starts gasma_greedy_batch on these arrays and returns a ticket; the arrays (inputs and `out`) must not be
        touched until greedy_wait(ticket) has returned."""
        n = len(read_off) - 1
        out, ops, op_off = self._async_out(n, want_ops, ops_cap, out)
        t = C.c_int(-1)
        self._check(self._L.gasma_greedy_submit(self._ctx, _np_ptr(reads), _np_ptr(read_off), _np_ptr(refs), _np_ptr(ref_off), n,
                                                _np_ptr(out["cost"]), _np_ptr(ops), 0 if ops is None else len(ops), _np_ptr(op_off),
                                                C.byref(t)))
        self._in_flight[t.value] = (out, n, want_ops, (reads, read_off, refs, ref_off))
        return t.value

    def greedy_submit_packed(self, r0, r1, read_off, f0, f1, ref_off, want_ops=True, ops_cap=None, out=None):
        n = len(read_off) - 1
        out, ops, op_off = self._async_out(n, want_ops, ops_cap, out)
        t = C.c_int(-1)
        self._check(self._L.gasma_greedy_submit_packed(self._ctx, _np_ptr(r0), _np_ptr(r1), _np_ptr(read_off), _np_ptr(f0), _np_ptr(f1),
                                                       _np_ptr(ref_off), n, _np_ptr(out["cost"]), _np_ptr(ops),
                                                       0 if ops is None else len(ops), _np_ptr(op_off), C.byref(t)))
        self._in_flight[t.value] = (out, n, want_ops, (r0, r1, read_off, f0, f1, ref_off))
        return t.value

    def greedy_wait(self, ticket):
        """Status of a submitted call (raises on error); returns its `out` dict."""
        rc = self._L.gasma_greedy_wait(self._ctx, ticket)
        out, n, want_ops, _keep = self._in_flight.pop(ticket, (None, 0, False, None))
        self._check(rc)
        if out is not None and want_ops:
            out["n_ops"] = int(out["op_off"][n])
        return out

    def leap_batch(self, reads, read_off, refs, ref_off):
        """LEAP comparator penalties (benchmark::_run_LEAP); INT32_MIN where no lane converges."""
        n = len(read_off) - 1
        pen = np.empty(n, np.int32)
        self._check(self._L.gasma_leap_batch(self._ctx, _np_ptr(reads), _np_ptr(read_off), _np_ptr(refs), _np_ptr(ref_off), n,
                                             _np_ptr(pen)))
        return pen

    def leap_batch_dev(self, d_reads, d_read_off, d_refs, d_ref_off, n, d_penalty):
        self._check(self._L.gasma_leap_batch_dev(self._ctx, _dev_ptr(d_reads), _dev_ptr(d_read_off), _dev_ptr(d_refs),
                                                 _dev_ptr(d_ref_off), n, _dev_ptr(d_penalty)))

    def nw_batch(self, reads, read_off, refs, ref_off, bound=None, want_ops=True, ops_cap=None):
        n = len(read_off) - 1
        out = {"penalty": np.empty(n, np.int32)}
        ops = op_off = None
        if want_ops:
            cap = ops_cap if ops_cap is not None else max(16, int(len(reads)) + int(len(refs)) + 2 * n)
            ops = out["ops"] = np.empty(cap, np.uint16)
            op_off = out["op_off"] = np.empty(n + 1, np.uint32)
        self._check(self._L.gasma_nw_batch(self._ctx, _np_ptr(reads), _np_ptr(read_off), _np_ptr(refs), _np_ptr(ref_off), n,
                                           _np_ptr(bound), _np_ptr(out["penalty"]), _np_ptr(ops),
                                           0 if ops is None else len(ops), _np_ptr(op_off)))
        if want_ops:
            out["n_ops"] = int(op_off[n])
        return out

    # -- device entry points (torch CUDA tensors) -----------------------------------------------
    def pack2bit_batch_dev(self, d_seqs, d_off, n, d_p0, d_p1):
        self._check(self._L.gasma_pack2bit_batch_dev(self._ctx, _dev_ptr(d_seqs), _dev_ptr(d_off), n, _dev_ptr(d_p0),
                                                     _dev_ptr(d_p1)))

    def greedy_batch_dev(self, d_reads, d_read_off, d_refs, d_ref_off, n, d_cost, d_ops=None, d_op_off=None):
        self._check(self._L.gasma_greedy_batch_dev(self._ctx, _dev_ptr(d_reads), _dev_ptr(d_read_off), _dev_ptr(d_refs),
                                                   _dev_ptr(d_ref_off), n, _dev_ptr(d_cost), _dev_ptr(d_ops),
                                                   0 if d_ops is None else d_ops.numel(), _dev_ptr(d_op_off)))

    def greedy_batch_dev_noalloc(self, d_reads, d_read_off, d_refs, d_ref_off, n, d_cost, d_ops=None, d_op_off=None):
        """The greedy step with its scratch in the library's static __device__ array: allocates nothing
        (n <= 8192 pairs, max_length <= 256; calls that overlap on one GPU race on the array)."""
        self._check(self._L.gasma_greedy_batch_dev_noalloc(self._ctx, _dev_ptr(d_reads), _dev_ptr(d_read_off), _dev_ptr(d_refs),
                                                           _dev_ptr(d_ref_off), n, _dev_ptr(d_cost), _dev_ptr(d_ops),
                                                           0 if d_ops is None else d_ops.numel(), _dev_ptr(d_op_off)))

    def nw_batch_dev(self, d_reads, d_read_off, d_refs, d_ref_off, n, d_bound, d_penalty, d_ops=None, d_op_off=None):
        self._check(self._L.gasma_nw_batch_dev(self._ctx, _dev_ptr(d_reads), _dev_ptr(d_read_off), _dev_ptr(d_refs),
                                               _dev_ptr(d_ref_off), n, _dev_ptr(d_bound), _dev_ptr(d_penalty),
                                               _dev_ptr(d_ops), 0 if d_ops is None else d_ops.numel(), _dev_ptr(d_op_off)))

    def metrics_dev(self, d_reads, d_read_off, n, g_cost, g_ops, g_off, nw_pen, nw_ops, nw_off, d_answers, d_counters):
        self._check(self._L.gasma_metrics_dev(self._ctx, _dev_ptr(d_reads), _dev_ptr(d_read_off), n, _dev_ptr(g_cost),
                                              _dev_ptr(g_ops), _dev_ptr(g_off), _dev_ptr(nw_pen), _dev_ptr(nw_ops),
                                              _dev_ptr(nw_off), _dev_ptr(d_answers), _dev_ptr(d_counters)))

    def generate_dev(self, cfg, first_pair, n, d_reads, d_read_off, d_refs, d_ref_off):
        self._check(self._L.gasma_generate_dev(self._ctx, C.byref(cfg), first_pair, n, _dev_ptr(d_reads),
                                               _dev_ptr(d_read_off), _dev_ptr(d_refs), _dev_ptr(d_ref_off)))

    def generate_ref_capacity(self, cfg, n):
        return int(self._L.gasma_generate_ref_capacity(C.byref(cfg), n))

    # -- the batched benchmark body (greedy + NW verifier + counters) -------------------------------
    def benchmark_batch(self, reads, read_off, refs, ref_off, answers=None, counters=None, want_penalties=False):
        """Host arrays in, the four counters of benchmark::_run_benchmark out (accumulated into `counters`)."""
        n = len(read_off) - 1
        if counters is None:
            counters = np.zeros(4, np.uint64)
        gc = np.empty(n, np.int32) if want_penalties else None
        nw = np.empty(n, np.int32) if want_penalties else None
        ms = (C.c_double * 2)(0.0, 0.0)
        self._check(self._L.gasma_benchmark_batch(self._ctx, _np_ptr(reads), _np_ptr(read_off), _np_ptr(refs),
                                                  _np_ptr(ref_off), n, _np_ptr(answers), _np_ptr(counters),
                                                  _np_ptr(gc), _np_ptr(nw), C.cast(ms, C.c_void_p)))
        return {"counters": counters, "greedy_cost": gc, "nw_penalty": nw, "device_ms": (ms[0], ms[1])}

    def benchmark_batch_dev(self, d_reads, d_read_off, d_refs, d_ref_off, n, d_counters, d_answers=None,
                            d_greedy_cost=None, d_nw_penalty=None):
        self._check(self._L.gasma_benchmark_batch_dev(self._ctx, _dev_ptr(d_reads), _dev_ptr(d_read_off), _dev_ptr(d_refs),
                                                      _dev_ptr(d_ref_off), n, _dev_ptr(d_answers), _dev_ptr(d_counters),
                                                      _dev_ptr(d_greedy_cost), _dev_ptr(d_nw_penalty)))

    def generate(self, cfg, first_pair, n):
        """Synthetic batch generated on the device, returned as host arrays (reads, read_off, refs, ref_off)."""
        reads = np.empty(n * cfg.length, np.uint8)
        refs = np.empty(max(1, self.generate_ref_capacity(cfg, n)), np.uint8)
        roff = np.empty(n + 1, np.uint32)
        foff = np.empty(n + 1, np.uint32)
        self._check(self._L.gasma_generate(self._ctx, C.byref(cfg), first_pair, n, _np_ptr(reads), _np_ptr(roff),
                                           _np_ptr(refs), _np_ptr(foff)))
        return reads, roff, refs[:int(foff[n])], foff
