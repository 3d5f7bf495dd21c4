"""ctypes helpers for tests/host_sim.cpp (the g++ build of the kernels' per-pair code)."""
import ctypes as C

import numpy as np


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def sim_greedy(S, ol, b, p, T=128, max_delta=None, mask_variant=0):
    """max_delta != None: the widened band (KParams::correction) with 2k+1+max_delta lane slots."""
    S.sim_greedy_batch.argtypes = [C.c_int] * 7 + [C.c_double] * 3 + [C.c_void_p] * 4 + [
        C.c_int64, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]
    S.sim_greedy_batch_corr.argtypes = [C.c_int] * 7 + [C.c_double] * 3 + [C.c_int] + [C.c_void_p] * 4 + [
        C.c_int64, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]
    stride = 2 * 64 * p.W + 8
    cost = np.zeros(b.n, np.int32)
    stage = np.zeros(b.n * stride, np.uint16)
    nops = np.zeros(b.n, np.uint32)
    st = np.zeros(1, np.uint32)
    steps = np.zeros(1, np.int64)
    head = (p.W, T, p.k, p.x, p.o, p.e, p.global_, p.match_prob, p.mismatch_prob, p.indel_prob)
    tail = (_p(b.reads), _p(b.read_off), _p(b.refs), _p(b.ref_off), b.n, _p(cost), _p(stage), stride, _p(nops), _p(st), _p(steps))
    if mask_variant:
        S.sim_greedy_batch_mask.argtypes = [C.c_int] + S.sim_greedy_batch.argtypes
        rc = S.sim_greedy_batch_mask(mask_variant, *head, *tail)
    elif max_delta is None:
        rc = S.sim_greedy_batch(*head, *tail)
    else:
        rc = S.sim_greedy_batch_corr(*head, max_delta, *tail)
    assert rc == 0
    text = [ol.render_ops(stage[i * stride:i * stride + nops[i]]) for i in range(b.n)]
    return cost, text, int(st[0]), int(steps[0])


def sim_pack(S, seqs, off, W, T=128):
    S.sim_pack_batch.argtypes = [C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p]
    n = len(off) - 1
    p0 = np.zeros((n, W), np.uint64)
    p1 = np.zeros((n, W), np.uint64)
    st = np.zeros(1, np.uint32)
    assert S.sim_pack_batch(W, T, _p(seqs), _p(off), n, _p(p0), _p(p1), _p(st)) == 0
    return p0, p1, int(st[0])


def sim_nw(S, b, W, x, o, e, bound=None, want_ops=True, cap_band=64, T=128):
    S.sim_nw_batch.argtypes = [C.c_int] * 5 + [C.c_void_p] * 4 + [C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int,
                                                                     C.c_void_p, C.c_int, C.c_void_p]
    stride = 2 * 64 * W + 8
    pen = np.zeros(b.n, np.int32)
    stage = np.zeros(b.n * stride, np.uint16) if want_ops else None
    nops = np.zeros(b.n, np.uint32)
    wide = np.zeros(1, np.int64)
    rc = S.sim_nw_batch(W, T, x, o, e, _p(b.reads), _p(b.read_off), _p(b.refs), _p(b.ref_off), b.n, _p(bound), _p(pen),
                        _p(stage), stride, _p(nops), cap_band, _p(wide))
    assert rc == 0
    ops = [stage[(i + 1) * stride - nops[i]:(i + 1) * stride] for i in range(b.n)] if want_ops else None
    return pen, ops, int(wide[0])


def sim_myers(S, b, W, want_ops=True, T=128, bound=None):
    S.sim_myers_batch.argtypes = [C.c_int] * 2 + [C.c_void_p] * 4 + [C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int,
                                                                    C.c_void_p, C.c_void_p]
    if bound is not None:
        bound = np.ascontiguousarray(bound, np.int32)
    stride = 2 * 64 * W + 8
    pen = np.zeros(b.n, np.int32)
    stage = np.zeros(b.n * stride, np.uint16) if want_ops else None
    nops = np.zeros(b.n, np.uint32)
    nwin = np.zeros(1, np.int64)
    rc = S.sim_myers_batch(W, T, _p(b.reads), _p(b.read_off), _p(b.refs), _p(b.ref_off), b.n, _p(bound), _p(pen), _p(stage),
                           stride, _p(nops), _p(nwin))
    assert rc == 0
    ops = [stage[(i + 1) * stride - nops[i]:(i + 1) * stride] for i in range(b.n)] if want_ops else None
    sim_myers.last_windowed = int(nwin[0])
    return pen, ops


def sim_counters(S, b, W, greedy, answers=None, T=128):
    """Fused verifier + counters of the benchmark body; returns (penalties, counters[4], deferred pairs)."""
    S.sim_counters_batch.argtypes = [C.c_int] * 2 + [C.c_void_p] * 4 + [C.c_int64] + [C.c_void_p] * 7
    pen = np.zeros(b.n, np.int32)
    counters = np.zeros(4, np.uint64)
    ndef = np.zeros(1, np.int64)
    gcost = np.ascontiguousarray(greedy["cost"], np.int32)
    gops = np.ascontiguousarray(greedy["ops"], np.uint16)
    goff = np.ascontiguousarray(greedy["op_off"], np.uint32)
    if answers is not None:
        answers = np.ascontiguousarray(answers, np.int32)
    rc = S.sim_counters_batch(W, T, _p(b.reads), _p(b.read_off), _p(b.refs), _p(b.ref_off), b.n, _p(gcost), _p(gops), _p(goff),
                              _p(answers), _p(pen), _p(counters), _p(ndef))
    assert rc == 0
    return pen, counters, int(ndef[0])


def sim_leap(S, b, W, k=3, x=1, o=1, e=1, T=128):
    S.sim_leap_batch.argtypes = [C.c_int] * 6 + [C.c_void_p] * 4 + [C.c_int64, C.c_void_p]
    pen = np.zeros(b.n, np.int32)
    rc = S.sim_leap_batch(W, T, k, x, o, e, _p(b.reads), _p(b.read_off), _p(b.refs), _p(b.ref_off), b.n, _p(pen))
    assert rc == 0
    return pen
