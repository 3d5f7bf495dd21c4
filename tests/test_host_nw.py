"""CPU tests of the NW / coverage / generator device code (g++ build, tests/host_sim.cpp) against the oracle."""
import ctypes as C

import numpy as np
import pytest

import sim_lib


@pytest.mark.parametrize("W", [1, 2, 3])
def test_nw_logic_matches_oracle(ol, host_sim, W):
    rng = np.random.default_rng(900 + W)
    for (x, o, e) in [(1, 1, 1), (2, 3, 1), (4, 6, 2), (1, 0, 1)]:
        b = ol.random_batch(rng, 400, 1, 64 * W, 0.3, 0.4, 50)
        r = ol.oracle_nw(b, x, o, e, threads=4)
        # (a) no bound: full matrix; (b) exact bound = optimum; (c) loose bound = greedy cost
        g = ol.oracle_greedy(b, ol.make_params(3, x, max(o, 1), e, W))["cost"] if o >= 1 else None
        for bound in (None, r["penalty"].astype(np.int32), g):
            if bound is None and o == 0:
                pass
            pen, ops, _ = sim_lib.sim_nw(host_sim, b, W, x, o, e, bound)
            assert pen.tolist() == r["penalty"].tolist(), (W, x, o, e, bound is None)
            exp = ol.render_all(r["ops"], r["op_off"])
            assert [ol.render_ops(v) for v in ops] == exp
    # the narrow/wide split of the kernels gives the same answers
    b = ol.random_batch(rng, 200, 1, 64 * W, 0.3, 0.4, 50)
    r = ol.oracle_nw(b, 1, 1, 1)
    pen, ops, wide = sim_lib.sim_nw(host_sim, b, W, 1, 1, 1, r["penalty"].astype(np.int32) + 9, cap_band=8)
    assert wide > 0 and pen.tolist() == r["penalty"].tolist()
    assert [ol.render_ops(v) for v in ops] == ol.render_all(r["ops"], r["op_off"])


@pytest.mark.parametrize("W", [1, 2, 3, 4])
def test_unit_penalty_bitvector_nw_matches_oracle(ol, host_sim, W):
    """The Myers-column verifier (x=o=e=1) gives the Gotoh oracle's penalty AND its traceback, op for op."""
    rng = np.random.default_rng(950 + W)
    for max_err, indel in ((0.05, 0.1), (0.3, 0.4), (0.9, 0.5)):
        b = ol.random_batch(rng, 600, 0, 64 * W, max_err, indel, 64 * W)
        r = ol.oracle_nw(b, 1, 1, 1, threads=4)
        pen, ops = sim_lib.sim_myers(host_sim, b, W)
        windowed = sim_lib.sim_myers.last_windowed
        assert pen.tolist() == r["penalty"].tolist()
        assert [ol.render_ops(v) for v in ops] == ol.render_all(r["ops"], r["op_off"])
        pen2, _ = sim_lib.sim_myers(host_sim, b, W, want_ops=False)
        assert pen2.tolist() == pen.tolist()
        # any upper bound on the optimum (the greedy penalty plays that role) gives the same penalty and the same walk:
        # exact, a little loose, and loose enough to push the band to 32 diagonals and beyond
        for slack in (0, 3, 14, 31, 40):
            bnd = r["penalty"].astype(np.int32) + rng.integers(0, slack + 1, b.n, dtype=np.int32)
            pen3, ops3 = sim_lib.sim_myers(host_sim, b, W, bound=bnd)
            assert pen3.tolist() == r["penalty"].tolist(), slack
            assert [ol.render_ops(v) for v in ops3] == ol.render_all(r["ops"], r["op_off"]), slack
            pen4, _ = sim_lib.sim_myers(host_sim, b, W, want_ops=False, bound=bnd)
            assert pen4.tolist() == r["penalty"].tolist(), slack
        # both traceback routes were exercised: 32-bit band windows and the wide (Gotoh) route
        if max_err <= 0.3:
            assert 0 < windowed
        if max_err >= 0.9 and W >= 2:
            assert windowed < b.n
    # unrelated strings of very different lengths (whole-matrix paths), and empty sides
    pairs = [(b"", b""), (b"ACGT", b""), (b"", b"ACGT"), (b"A" * (64 * W), b"T" * (64 * W)), (b"ACGT" * (16 * W), b"TGCA" * 5)]
    b = ol.Batch.from_pairs(pairs)
    r = ol.oracle_nw(b, 1, 1, 1)
    pen, ops = sim_lib.sim_myers(host_sim, b, W)
    assert pen.tolist() == r["penalty"].tolist()
    assert [ol.render_ops(v) for v in ops] == ol.render_all(r["ops"], r["op_off"])


def test_coverage_logic_matches_oracle(ol, host_sim):
    host_sim.sim_coverage.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_void_p, C.c_int]
    rng = np.random.default_rng(31)
    b = ol.random_batch(rng, 1500, 20, 128, 0.25, 0.3, 50)
    g = ol.oracle_greedy(b, ol.make_params())
    nw = ol.oracle_nw(b, 1, 1, 1)
    exp = ol.oracle_metrics(b, g, nw)
    cov = 0
    for i in range(b.n):
        s1 = b.reads[b.read_off[i]:b.read_off[i + 1]]
        go = g["ops"][g["op_off"][i]:g["op_off"][i + 1]]
        no = nw["ops"][nw["op_off"][i]:nw["op_off"][i + 1]]
        cov += host_sim.sim_coverage(sim_lib._p(s1), len(s1), sim_lib._p(go), len(go), sim_lib._p(no), len(no))
    assert exp[0] == b.n and exp[1] == b.n
    assert exp[2] == int((g["cost"] == nw["penalty"]).sum())
    assert cov == int(exp[3]) and 0 < cov < b.n


@pytest.mark.parametrize("W", [1, 2, 3, 4])
def test_fused_verifier_counters_match_oracle(ol, host_sim, W):
    """nw_band_kernel's counter mode: no NW CIGAR, the walk's '=' runs feed the coverage test on 2-bit planes."""
    rng = np.random.default_rng(4100 + W)
    seen_deferred = 0
    for max_err, indel, k in ((0.08, 0.1, 3), (0.25, 0.3, 3), (0.25, 0.04, 5), (0.7, 0.4, 3)):
        b = ol.random_batch(rng, 700, 0, 64 * W, max_err, indel, 50)
        g = ol.oracle_greedy(b, ol.make_params(k, 1, 1, 1, W, 1))
        nw = ol.oracle_nw(b, 1, 1, 1, threads=4)
        exp = ol.oracle_metrics(b, g, nw)
        pen, counters, ndef = sim_lib.sim_counters(host_sim, b, W, g)
        seen_deferred += ndef
        assert pen.tolist() == nw["penalty"].tolist()
        assert counters.tolist() == exp.tolist()
        assert 0 < counters[3] < b.n or max_err < 0.1
        # stored answers (read_answer_file) replace the NW penalty as the truth for some pairs
        ans = np.where(rng.random(b.n) < 0.5, nw["penalty"] + rng.integers(0, 2, b.n), np.iinfo(np.int32).min).astype(np.int32)
        exp2 = ol.oracle_metrics(b, g, nw, ans)
        _, c2, _ = sim_lib.sim_counters(host_sim, b, W, g, answers=ans)
        assert c2.tolist() == exp2.tolist()
    if W >= 2:
        assert seen_deferred > 0     # the Gotoh + character-based route was exercised too


@pytest.mark.parametrize("W", [1, 2, 3, 4])
def test_leap_logic_matches_oracle(ol, host_sim, W):
    """leap_pair (lane bit masks + ring tables) == the LV restatement, which is pinned to the unmodified class."""
    rng = np.random.default_rng(5200 + W)
    for (k, x, o, e) in ((3, 1, 1, 1), (5, 1, 2, 1), (3, 2, 3, 1), (8, 4, 6, 2)):
        for max_err, indel in ((0.08, 0.2), (0.3, 0.4), (0.95, 0.5)):
            b = ol.random_batch(rng, 500, 0, 64 * W, max_err, indel, 50)
            exp = ol.oracle_leap(b, k, x, o, e, threads=4)
            got = sim_lib.sim_leap(host_sim, b, W, k, x, o, e)
            assert got.tolist() == exp.tolist(), (k, x, o, e, max_err)
    pairs = [(b"", b""), (b"ACGT", b""), (b"", b"ACGT"), (b"A" * (64 * W), b"T" * (64 * W)), (b"ACGT" * (16 * W), b"ACGT" * (16 * W)),
             (b"ACGTACGTAC", b"ACGTTACGTAC"), (b"C" * 40, b"C" * 37)]
    b = ol.Batch.from_pairs(pairs)
    assert sim_lib.sim_leap(host_sim, b, W).tolist() == ol.oracle_leap(b).tolist()


def test_generator_logic_matches_oracle_twin(ol, host_sim):
    host_sim.sim_generate_pair.argtypes = [C.c_uint64] + [C.c_int] * 3 + [C.c_uint64] * 5 + [C.c_void_p] * 3
    for mode, e, prof in [(0, 0.10, None), (1, 0.20, None), (2, 0.0, (0.02452, 0.0468, 0.0553)), (0, 0.15, None)]:
        cfg = ol.gen_config(77, 100, e, mode=mode, profile=prof)
        ref = ol.generate(cfg, 500, 300)
        for i in range(300):
            rd = np.zeros(128, np.uint8); rf = np.zeros(256, np.uint8); ln = C.c_int(0)
            n = host_sim.sim_generate_pair(cfg.seed, cfg.length, cfg.mode, cfg.num_errors, cfg.thr_mismatch, cfg.thr_sub,
                                           cfg.thr_ins, cfg.thr_del, 500 + i, sim_lib._p(rd), sim_lib._p(rf), C.byref(ln))
            a, f = ref.pair(i)
            assert n == ln.value == len(f)
            assert rd[:100].tobytes() == a and rf[:n].tobytes() == f
