"""ctypes bindings for the checker libraries under oracle/ (TEST INFRASTRUCTURE).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs
import this module.  The product never does.

  * oracle/libgasma_oracle.so       plain-C restatement (always buildable)
  * oracle/_ref/libgasma_ref.so     UNMODIFIED reference compiled in place (needs /root/reference at
                                    build time; the prebuilt .so travels to the GPU box)
"""
import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")
OP_LETTERS = "MID=X"
DEFAULT_PROBS = (0.80, 0.20 / 3, 0.40 / 3)  # hurdle_matrix.h:557-559


def build_oracle(quiet=True):
    """(Re)build the checker libs; never fails just because /root/reference is absent."""
    subprocess.run(["make", "-C", ORACLE_DIR], check=True,
                   stdout=subprocess.DEVNULL if quiet else None)


class GoParams(C.Structure):
    _fields_ = [("k", C.c_int), ("x", C.c_int), ("o", C.c_int), ("e", C.c_int),
                ("global_", C.c_int), ("W", C.c_int),
                ("match_prob", C.c_double), ("mismatch_prob", C.c_double), ("indel_prob", C.c_double),
                ("correction", C.c_int), ("mask_variant", C.c_int)]


class GoGenConfig(C.Structure):
    _fields_ = [("seed", C.c_uint64), ("length", C.c_int), ("mode", C.c_int), ("num_errors", C.c_int),
                ("thr_mismatch", C.c_uint64), ("thr_sub", C.c_uint64), ("thr_ins", C.c_uint64),
                ("thr_del", C.c_uint64)]


def _ptr(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


_oracle = None
_ref = None
_ref_corr = None


def oracle():
    global _oracle
    if _oracle is None:
        path = os.path.join(ORACLE_DIR, "libgasma_oracle.so")
        if not os.path.exists(path):
            build_oracle()
        L = C.CDLL(path)
        L.go_greedy_batch.restype = C.c_int
        L.go_greedy_batch.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64,
                                      C.POINTER(GoParams), C.c_int, C.c_void_p, C.c_void_p, C.c_int,
                                      C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p]
        L.go_nw_batch.restype = C.c_int
        L.go_nw_batch.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64,
                                  C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p,
                                  C.c_void_p, C.c_void_p, C.c_int64]
        L.go_pack2bit.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p]
        L.go_metrics.argtypes = [C.c_void_p, C.c_void_p, C.c_int64] + [C.c_void_p] * 8
        L.go_generate_pair.restype = C.c_int
        L.go_generate_pair.argtypes = [C.POINTER(GoGenConfig), C.c_uint64, C.c_void_p, C.c_void_p]
        L.go_gen_ref_capacity.restype = C.c_int
        L.go_gen_ref_capacity.argtypes = [C.POINTER(GoGenConfig)]
        L.go_lcm.restype = C.c_int
        L.go_lcm.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_int]
        L.go_covers.restype = C.c_int
        L.go_covers.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_int]
        L.go_leap_batch.restype = C.c_int
        L.go_leap_batch.argtypes = [C.c_void_p] * 4 + [C.c_int64] + [C.c_int] * 5 + [C.c_void_p]
        _oracle = L
    return _oracle


_fast = None


def fast():
    """oracle/libgasma_fast.so: the fast CPU arm (W fixed at compile time), bench.py's CPU baseline beyond 128 characters."""
    global _fast
    if _fast is None:
        path = os.path.join(ORACLE_DIR, "libgasma_fast.so")
        if not os.path.exists(path):
            build_oracle()
        L = C.CDLL(path)
        L.gf_greedy_batch.restype = C.c_int
        L.gf_greedy_batch.argtypes = ([C.c_void_p] * 4 + [C.c_int64] + [C.c_int] * 6 + [C.c_double] * 3 + [C.c_int] +
                                      [C.c_void_p] * 3 + [C.c_int64, C.c_void_p])
        _fast = L
    return _fast


def fast_greedy(b, k=3, x=1, o=1, e=1, W=2, global_=1, probs=DEFAULT_PROBS, threads=1, want_ops=True, ops_cap=None):
    """cost + CIGAR CSR from the fast CPU arm (one pass; want_ops=False still builds every CIGAR, like get_CIGAR())."""
    cost = np.zeros(b.n, np.int32)
    op_off = np.zeros(b.n + 1, np.uint32)
    cap = ops_cap if ops_cap is not None else max(16, 24 * b.n)
    ops = np.zeros(cap, np.uint16) if want_ops else None
    steps = C.c_int64(0)
    rc = fast().gf_greedy_batch(_ptr(b.reads), _ptr(b.read_off), _ptr(b.refs), _ptr(b.ref_off), b.n, k, x, o, e, global_, W,
                                probs[0], probs[1], probs[2], threads, _ptr(cost), _ptr(ops), _ptr(op_off), cap, C.byref(steps))
    if rc == -4 and ops_cap is None:
        return fast_greedy(b, k, x, o, e, W, global_, probs, threads, True, int(op_off[-1]))
    assert rc == 0, rc
    out = {"cost": cost, "op_off": op_off, "steps": steps.value}
    if want_ops:
        out["ops"] = ops[:int(op_off[-1])]
    return out


def ref_available():
    return os.path.exists(os.path.join(ORACLE_DIR, "_ref", "libgasma_ref.so"))


def ref():
    global _ref
    if _ref is None:
        L = C.CDLL(os.path.join(ORACLE_DIR, "_ref", "libgasma_ref.so"))
        L.gasma_ref_greedy_batch.restype = C.c_int
        L.gasma_ref_greedy_batch.argtypes = ([C.c_void_p] * 4 + [C.c_int64] + [C.c_int] * 4 + [C.c_double] * 3 +
                                             [C.c_int] * 2 + [C.c_void_p, C.c_void_p, C.c_int])
        L.gasma_ref_convert2bit1.argtypes = [C.c_void_p] * 4
        L.gasma_ref_lcm.restype = C.c_int
        L.gasma_ref_lcm.argtypes = [C.c_char_p, C.c_char_p, C.c_char_p, C.c_int, C.c_void_p, C.c_int]
        L.gasma_ref_covers.restype = C.c_int
        L.gasma_ref_covers.argtypes = [C.c_char_p, C.c_char_p]
        L.gasma_ref_significance.argtypes = [C.c_double] * 3 + [C.c_void_p]
        if hasattr(L, "gasma_ref_leap_batch"):
            L.gasma_ref_leap_batch.restype = C.c_int
            L.gasma_ref_leap_batch.argtypes = [C.c_void_p] * 4 + [C.c_int64] + [C.c_int] * 6 + [C.c_void_p] * 2
        _ref = L
    return _ref


def ref_corr_available():
    return os.path.exists(os.path.join(ORACLE_DIR, "_ref", "libgasma_ref_corr.so"))


def ref_corr():
    """The unmodified reference compiled with -DCORRECTION (hurdle_matrix.h:501-512, :634-641): a library of its
    own, because the macro changes the one hurdle_matrix<int_128bit> instantiation."""
    global _ref_corr
    if _ref_corr is None:
        L = C.CDLL(os.path.join(ORACLE_DIR, "_ref", "libgasma_ref_corr.so"))
        L.gasma_ref_greedy_batch.restype = C.c_int
        L.gasma_ref_greedy_batch.argtypes = ([C.c_void_p] * 4 + [C.c_int64] + [C.c_int] * 4 + [C.c_double] * 3 +
                                             [C.c_int] * 2 + [C.c_void_p, C.c_void_p, C.c_int])
        _ref_corr = L
    return _ref_corr


_ref_mask = {}


def ref_mask_available(variant):
    return os.path.exists(os.path.join(ORACLE_DIR, "_ref", "libgasma_ref_mask%d.so" % variant))


def ref_mask(variant):
    """The unmodified reference with the highway search on another mask variant (hurdle_matrix.h:453, utils.h:200-240):
    1 = flip_short_hurdles(2), 2 = flip_short_hurdles(1).flip_short_matches(1).  See oracle/ref_driver.cpp."""
    if variant not in _ref_mask:
        L = C.CDLL(os.path.join(ORACLE_DIR, "_ref", "libgasma_ref_mask%d.so" % variant))
        L.gasma_ref_greedy_batch.restype = C.c_int
        L.gasma_ref_greedy_batch.argtypes = ([C.c_void_p] * 4 + [C.c_int64] + [C.c_int] * 4 + [C.c_double] * 3 +
                                             [C.c_int] * 2 + [C.c_void_p, C.c_void_p, C.c_int])
        _ref_mask[variant] = L
    return _ref_mask[variant]


# ----------------------------------------------------------------------------------------------
# batch containers: concatenated ASCII + CSR offsets (the layout the C ABI uses)
# ----------------------------------------------------------------------------------------------
class Batch:
    def __init__(self, reads, read_off, refs, ref_off):
        self.reads = np.ascontiguousarray(reads, dtype=np.uint8)
        self.refs = np.ascontiguousarray(refs, dtype=np.uint8)
        self.read_off = np.ascontiguousarray(read_off, dtype=np.uint32)
        self.ref_off = np.ascontiguousarray(ref_off, dtype=np.uint32)
        self.n = len(self.read_off) - 1

    @staticmethod
    def from_pairs(pairs):
        rl = [len(p[0]) for p in pairs]
        fl = [len(p[1]) for p in pairs]
        ro = np.zeros(len(pairs) + 1, np.uint32)
        fo = np.zeros(len(pairs) + 1, np.uint32)
        ro[1:] = np.cumsum(rl)
        fo[1:] = np.cumsum(fl)
        reads = np.frombuffer(b"".join(p[0] for p in pairs), dtype=np.uint8)
        refs = np.frombuffer(b"".join(p[1] for p in pairs), dtype=np.uint8)
        return Batch(reads, ro, refs, fo)

    def pair(self, i):
        return (self.reads[self.read_off[i]:self.read_off[i + 1]].tobytes(),
                self.refs[self.ref_off[i]:self.ref_off[i + 1]].tobytes())

    def slice(self, lo, hi):
        ro = self.read_off[lo:hi + 1].astype(np.int64)
        fo = self.ref_off[lo:hi + 1].astype(np.int64)
        return Batch(self.reads[ro[0]:ro[-1]], ro - ro[0], self.refs[fo[0]:fo[-1]], fo - fo[0])


def batch_from_strings(reads, refs):
    """Batch from two lists of str (e.g. a parsed '>read / <ref' file)."""
    return Batch.from_pairs([(r.encode(), f.encode()) for r, f in zip(reads, refs)])


def make_params(k=3, x=1, o=1, e=1, W=2, global_=1, probs=DEFAULT_PROBS, correction=0, mask_variant=0):
    return GoParams(k, x, o, e, global_, W, probs[0], probs[1], probs[2], correction, mask_variant)


def gen_config(seed, length, error_rate, mismatch_rate=0.96, mode=0, profile=None):
    """mode 0: exact ceil(L*e) edits (benchmark.cpp:19 passes exact=true), 1: uniform, 2: per-base profile
    (profile = (p_sub, p_ins, p_del))."""
    ne = int(np.ceil(np.float32(length) * np.float32(error_rate)))  # float math as benchmark_dataset.h:153
    c = GoGenConfig(seed, length, mode, ne, int(mismatch_rate * 4294967296.0), 0, 0, 0)
    if mode == 2:
        ps, pi, pd = profile
        c.thr_del = int(pd * 4294967296.0)
        c.thr_ins = c.thr_del + int(pi * 4294967296.0)
        c.thr_sub = c.thr_ins + int(ps * 4294967296.0)
    return c


def generate(cfg, first, n):
    """CPU twin of the device generator: pairs first .. first+n."""
    L = oracle()
    cap = L.go_gen_ref_capacity(C.byref(cfg))
    reads = np.zeros(n * cfg.length, np.uint8)
    refs = np.zeros(n * cap, np.uint8)
    ro = (np.arange(n + 1, dtype=np.uint64) * cfg.length).astype(np.uint32)
    fo = np.zeros(n + 1, np.uint32)
    tmp = np.zeros(cap + 16, np.uint8)
    pos = 0
    for i in range(n):
        ln = L.go_generate_pair(C.byref(cfg), first + i, _ptr(reads[i * cfg.length:]), _ptr(tmp))
        refs[pos:pos + ln] = tmp[:ln]
        pos += ln
        fo[i + 1] = pos
    return Batch(reads, ro, refs[:pos].copy(), fo)


def random_batch(rng, n, min_len=1, max_len=128, max_err=0.3, indel_share=0.3, max_diff=50, alphabet=b"ACGT"):
    """numpy-driven random pairs for oracle-vs-reference fuzzing (lengths ragged, |n-m| <= max_diff)."""
    pairs = []
    alpha = np.frombuffer(alphabet, dtype=np.uint8)
    for _ in range(n):
        m = int(rng.integers(min_len, max_len + 1))
        read = alpha[rng.integers(0, len(alpha), m)]
        ref = list(read)
        ne = int(rng.integers(0, int(m * max_err) + 1))
        for _ in range(ne):
            r = rng.random()
            if r >= indel_share or not ref:
                if ref:
                    ref[int(rng.integers(0, len(ref)))] = alpha[rng.integers(0, len(alpha))]
            elif rng.random() < 0.5:
                del ref[int(rng.integers(0, len(ref)))]
            else:
                ref.insert(int(rng.integers(0, len(ref) + 1)), alpha[rng.integers(0, len(alpha))])
        ref = ref[:max_len]
        while abs(len(ref) - m) > max_diff:
            ref = ref[:-1] if len(ref) > m else ref + [alpha[0]]
        pairs.append((bytes(bytearray(read.tolist())), bytes(bytearray(int(c) for c in ref))))
    return Batch.from_pairs(pairs)


# ----------------------------------------------------------------------------------------------
# oracle calls
# ----------------------------------------------------------------------------------------------
def oracle_greedy(b, params, threads=1, want_text=False, stride=2048):
    L = oracle()
    cost = np.zeros(b.n, np.int32)
    op_off = np.zeros(b.n + 1, np.uint32)
    text = np.zeros(b.n * stride, np.uint8) if want_text else None
    steps = C.c_int64(0)
    rc = L.go_greedy_batch(_ptr(b.reads), _ptr(b.read_off), _ptr(b.refs), _ptr(b.ref_off), b.n,
                           C.byref(params), threads, _ptr(cost), _ptr(text), stride,
                           None, _ptr(op_off), 0, C.byref(steps))
    assert rc == 0, rc
    ops = np.zeros(max(int(op_off[-1]), 1), np.uint16)
    rc = L.go_greedy_batch(_ptr(b.reads), _ptr(b.read_off), _ptr(b.refs), _ptr(b.ref_off), b.n,
                           C.byref(params), threads, None, None, 0,
                           _ptr(ops), _ptr(op_off), len(ops), None)
    assert rc == 0, rc
    out = {"cost": cost, "ops": ops[:int(op_off[-1])], "op_off": op_off, "steps": steps.value}
    if want_text:
        out["text"] = [text[i * stride:(i + 1) * stride].tobytes().split(b"\0")[0].decode() for i in range(b.n)]
    return out


def oracle_nw(b, x=1, o=1, e=1, threads=1, want_ops=True):
    L = oracle()
    pen = np.zeros(b.n, np.int32)
    if not want_ops:
        L.go_nw_batch(_ptr(b.reads), _ptr(b.read_off), _ptr(b.refs), _ptr(b.ref_off), b.n, x, o, e, threads,
                      _ptr(pen), None, None, 0)
        return {"penalty": pen}
    op_off = np.zeros(b.n + 1, np.uint32)
    L.go_nw_batch(_ptr(b.reads), _ptr(b.read_off), _ptr(b.refs), _ptr(b.ref_off), b.n, x, o, e, threads,
                  _ptr(pen), None, _ptr(op_off), 0)
    ops = np.zeros(max(int(op_off[-1]), 1), np.uint16)
    L.go_nw_batch(_ptr(b.reads), _ptr(b.read_off), _ptr(b.refs), _ptr(b.ref_off), b.n, x, o, e, threads,
                  None, _ptr(ops), _ptr(op_off), len(ops))
    return {"penalty": pen, "ops": ops[:int(op_off[-1])], "op_off": op_off}


def oracle_metrics(b, greedy, nw, answers=None):
    out = np.zeros(4, np.uint64)
    oracle().go_metrics(_ptr(b.reads), _ptr(b.read_off), b.n, _ptr(greedy["cost"]), _ptr(greedy["ops"]),
                        _ptr(greedy["op_off"]), _ptr(nw["penalty"]), _ptr(nw["ops"]), _ptr(nw["op_off"]),
                        _ptr(answers), _ptr(out))
    return out


def oracle_pack(seq, W):
    p0 = np.zeros(W, np.uint64)
    p1 = np.zeros(W, np.uint64)
    s = np.frombuffer(seq, dtype=np.uint8)
    oracle().go_pack2bit(_ptr(s) if len(s) else None, len(s), W, _ptr(p0), _ptr(p1))
    return p0, p1


def render_ops(ops):
    return "".join("%d%s" % (int(o) >> 3, OP_LETTERS[int(o) & 7]) for o in ops)


def render_all(ops, op_off):
    return [render_ops(ops[op_off[i]:op_off[i + 1]]) for i in range(len(op_off) - 1)]


# ----------------------------------------------------------------------------------------------
# unmodified reference calls
# ----------------------------------------------------------------------------------------------
def ref_greedy(b, k=3, x=1, o=1, e=1, probs=DEFAULT_PROBS, mode=0, threads=1, want_text=True, stride=1280,
               correction=False, mask_variant=0):
    L = ref_corr() if correction else (ref_mask(mask_variant) if mask_variant else ref())
    cost = np.zeros(b.n, np.int32)
    text = np.zeros(b.n * stride, np.uint8) if want_text else None
    need = L.gasma_ref_greedy_batch(_ptr(b.reads), _ptr(b.read_off), _ptr(b.refs), _ptr(b.ref_off), b.n,
                                    k, x, o, e, probs[0], probs[1], probs[2], mode, threads,
                                    _ptr(cost), _ptr(text), stride)
    assert need <= stride, "CIGAR truncated"
    out = {"cost": cost}
    if want_text:
        out["text"] = [text[i * stride:(i + 1) * stride].tobytes().split(b"\0")[0].decode() for i in range(b.n)]
    return out


def oracle_leap(b, k=3, x=1, o=1, e=1, threads=1):
    """LEAP comparator penalties (plain-C restatement); INT32_MIN = no lane converged."""
    pen = np.zeros(b.n, np.int32)
    rc = oracle().go_leap_batch(_ptr(b.reads), _ptr(b.read_off), _ptr(b.refs), _ptr(b.ref_off), b.n, k, x, o, e, threads, _ptr(pen))
    assert rc == 0
    return pen


def ref_leap(b, k=3, x=1, o=1, e=1, mode=0, threads=1):
    """The unmodified class LV driven like benchmark::_run_LEAP; mode 0 = one object per thread, 1 = fresh per pair."""
    pen = np.zeros(b.n, np.int32)
    ok = np.zeros(b.n, np.uint8)
    rc = ref().gasma_ref_leap_batch(_ptr(b.reads), _ptr(b.read_off), _ptr(b.refs), _ptr(b.ref_off), b.n, k, x, o, e, mode,
                                    threads, _ptr(pen), _ptr(ok))
    assert rc == 0
    return pen, ok


def ref_convert2bit1(buf128):
    a = np.frombuffer(buf128, dtype=np.uint8).copy()
    assert len(a) == 128
    b0 = np.zeros(16, np.uint8)
    b1 = np.zeros(16, np.uint8)
    sh = np.zeros(128, np.uint8)
    ref().gasma_ref_convert2bit1(_ptr(a), _ptr(b0), _ptr(b1), _ptr(sh))
    return b0, b1, sh
