"""CPU tests: the plain-C oracle against the golden vectors captured from the unmodified reference
(tests/golden/, made by tests/golden/make_golden.py) and, where oracle/_ref was built, against the
reference itself on fresh seeded batches."""
import numpy as np
import pytest


def _batch(ol, pairs):
    return ol.Batch.from_pairs([(a.encode("latin1"), b.encode("latin1")) for a, b in pairs])


def test_oracle_matches_golden_greedy(ol, golden):
    g = golden("greedy_ref.json")
    total = 0
    for s in g["sets"]:
        ps = s["params"]
        p = ol.make_params(ps["k"], ps["x"], ps["o"], ps["e"], 2, 1, tuple(ps["probs"]))
        b = _batch(ol, s["pairs"])
        r = ol.oracle_greedy(b, p, threads=4, want_text=True)
        assert r["cost"].tolist() == s["cost"], s["name"]
        assert r["text"] == s["cigar"], s["name"]
        assert ol.render_all(r["ops"], r["op_off"]) == s["cigar"]
        total += b.n
    assert total > 9000


def test_golden_headline_vectors(golden):
    """G1..G8 of SURVEY.md section 8c, literally."""
    s = golden("greedy_ref.json")["sets"][0]
    assert s["params"]["k"] == 3
    got = list(zip(s["cost"], s["cigar"]))
    assert got[0] == (6, "22M1D50M1D28M")
    assert got[1] == (47, "5M1I19M1I12M1I2M2D3M2I1M6M1D3M1D3M1D1M2I4M2D8M1I4M2I1M2D1M2I3M8M1I")
    assert got[2] == (6, "7M1I44M1D31M1D17M")
    assert got[3:8] == [(0, "4M"), (1, "1M1I2M"), (4, "4M"), (0, ""), (0, "128M")]


def test_oracle_pack_matches_golden(ol, golden):
    for v in golden("convert2bit_ref.json")["vectors"]:
        p0, p1 = ol.oracle_pack(v["in"].encode("latin1"), 2)
        assert p0.view(np.uint8).tolist() == v["bits0"]
        assert p1.view(np.uint8).tolist() == v["bits1"]


def test_significance_constants(golden):
    import math
    sig = [float.fromhex(h) for h in golden("convert2bit_ref.json")["significance_default"]]
    assert sig[0] == math.log(0.80 / 0.25)
    assert sig[1] == sig[2]  # the tie the FP64 heuristic hinges on (SURVEY.md hazard 2)


def test_coverage_matches_golden(ol, golden):
    import ctypes as C
    g = golden("coverage_ref.json")
    codes = {"M": 0, "I": 1, "D": 2, "=": 3, "X": 4}
    import re
    for v in g["lcm"]:
        ops = np.array([(int(n) << 3) | codes[c] for n, c in re.findall(r"(\d+)([MIDX=])", v["cigar"])], np.uint16)
        s1 = np.frombuffer(v["s1"].encode(), np.uint8)
        out = np.zeros(600, np.uint8)
        n = ol.oracle().go_lcm(ol._ptr(s1), len(s1), ol._ptr(ops), len(ops), v["threshold"], ol._ptr(out), 600)
        assert out[:n].tobytes().decode() == v["lcm"]
    for v in g["covers"]:
        a = np.frombuffer(v["a"].encode() + b"\0", np.uint8)
        b = np.frombuffer(v["b"].encode() + b"\0", np.uint8)
        assert ol.oracle().go_covers(ol._ptr(a), len(v["a"]), ol._ptr(b), len(v["b"])) == v["covers"]


def test_oracle_vs_reference_fresh(ol):
    """Fresh seeded fuzz against the unmodified reference (only where oracle/_ref could be built)."""
    if not ol.ref_available():
        pytest.skip("oracle/_ref not built (needs /root/reference)")
    rng = np.random.default_rng(4242)
    for (k, x, o, e) in [(3, 1, 1, 1), (6, 2, 3, 1)]:
        b = ol.random_batch(rng, 4000, 1, 128, 0.3, 0.3, 50)
        r = ol.oracle_greedy(b, ol.make_params(k, x, o, e), threads=4, want_text=True)
        f = ol.ref_greedy(b, k, x, o, e, threads=4)
        assert r["cost"].tolist() == f["cost"].tolist()
        assert r["text"] == f["text"]


def test_oracle_semi_global_vs_reference(ol):
    """SEMI_GLOBAL (no switch cost on the first step, no final lane switch; hurdle_matrix.h:314,336,578) against the
    unmodified reference constructed with that alignment type."""
    if not ol.ref_available():
        pytest.skip("oracle/_ref not built (needs /root/reference)")
    rng = np.random.default_rng(4343)
    for (k, x, o, e) in [(3, 1, 1, 1), (5, 2, 3, 1)]:
        b = ol.random_batch(rng, 3000, 1, 128, 0.3, 0.3, 50)
        r = ol.oracle_greedy(b, ol.make_params(k, x, o, e, 2, 0), threads=4, want_text=True)
        f = ol.ref_greedy(b, k, x, o, e, mode=2, threads=4)
        assert r["cost"].tolist() == f["cost"].tolist()
        assert r["text"] == f["text"]
        glob = ol.ref_greedy(b, k, x, o, e, mode=0, threads=4)
        assert (f["cost"] != glob["cost"]).mean() > 0.05     # the mode matters on this data


def test_generator_is_deterministic_and_shaped(ol):
    cfg = ol.gen_config(0x47415341, 100, 0.10)
    assert cfg.num_errors == 10
    assert ol.gen_config(1, 100, 0.15).num_errors == 16  # float ceil as benchmark_dataset.h:153 (100*0.15f > 15)
    a = ol.generate(cfg, 1000, 50)
    b = ol.generate(cfg, 1010, 10)
    assert a.slice(10, 20).reads.tobytes() == b.reads.tobytes()
    assert a.slice(10, 20).refs.tobytes() == b.refs.tobytes()
    lens = np.diff(a.ref_off.astype(np.int64))
    assert lens.min() >= 90 and lens.max() <= 110
    assert set(a.reads.tolist()) <= set(b"ACGT")


def test_nw_oracle_basics(ol):
    b = ol.Batch.from_pairs([(b"ACGT", b"ACGT"), (b"ACGT", b"AGT"), (b"AAAA", b"TTTT"), (b"", b"ACG"), (b"ACGTACGT", b"ACGACGTT")])
    r = ol.oracle_nw(b, 1, 1, 1)
    assert r["penalty"].tolist() == [0, 1, 4, 3, 2]
    assert ol.render_all(r["ops"], r["op_off"])[0] == "4="
    r2 = ol.oracle_nw(b, 2, 3, 1)
    assert r2["penalty"].tolist() == [0, 3, 8, 5, 6]
    # every traceback is consistent: it consumes both strings and re-scores to the penalty
    for i, text in enumerate(ol.render_all(r2["ops"], r2["op_off"])):
        import re
        pen = ri = fi = 0
        for n, c in re.findall(r"(\d+)([=XID])", text):
            n = int(n)
            if c == "=": ri += n; fi += n
            elif c == "X": ri += n; fi += n; pen += 2 * n
            elif c == "I": ri += n; pen += 3 + (n - 1)
            else: fi += n; pen += 3 + (n - 1)
        a, f = b.pair(i)
        assert (ri, fi, pen) == (len(a), len(f), int(r2["penalty"][i]))


def test_oracle_leap_matches_golden(ol, golden):
    """LEAP comparator (class LV, LV_BAG.cpp) restatement against penalties captured from the unmodified class."""
    g = golden("greedy_ref.json")
    lp = golden("leap_ref.json")
    total = 0
    for s in lp["sets"]:
        b = _batch(ol, g["sets"][s["set_index"]]["pairs"])
        pen = ol.oracle_leap(b, s["k"], s["x"], s["o"], s["e"], threads=4)
        assert pen.tolist() == s["penalty"], (s["name"], s["k"], s["x"], s["o"], s["e"])
        total += b.n
    assert total > 10000
    # benchmark setting (k=3, x=o=e=1) on the benchmark's own kind of data: LEAP is optimal far more often than not
    first = next(s for s in lp["sets"] if s["name"] == "gen_e05" and s["k"] == 3 and s["x"] == 1)
    assert min(first["penalty"]) >= 0 and max(first["penalty"]) <= 5


def test_oracle_leap_vs_reference_fresh(ol):
    """Fresh fuzz against the unmodified LV class, both reused across pairs (what the benchmark does) and
    constructed per pair: the results do not depend on earlier pairs."""
    if not ol.ref_available() or not hasattr(ol.ref(), "gasma_ref_leap_batch"):
        pytest.skip("oracle/_ref not built (needs /root/reference)")
    rng = np.random.default_rng(77)
    for (k, x, o, e) in [(3, 1, 1, 1), (4, 1, 2, 1), (8, 4, 6, 2)]:
        for max_err in (0.1, 0.9):
            b = ol.random_batch(rng, 2500, 0, 128, max_err, 0.4, 50)
            a = ol.oracle_leap(b, k, x, o, e, threads=4)
            r0, _ = ol.ref_leap(b, k, x, o, e, mode=0, threads=1)
            r1, _ = ol.ref_leap(b, k, x, o, e, mode=1, threads=4)
            assert a.tolist() == r0.tolist() == r1.tolist()


def test_oracle_correction_matches_golden(ol, golden):
    """The widened band of the reference's -DCORRECTION build (hurdle_matrix.h:501-512, :634-641) against the golden
    results of the unmodified reference compiled with that macro (tests/golden/make_golden_correction.py)."""
    total = 0
    for s in golden("greedy_ref_correction.json")["sets"]:
        ps = s["params"]
        p = ol.make_params(ps["k"], ps["x"], ps["o"], ps["e"], 2, 1, tuple(ps["probs"]), correction=1)
        b = _batch(ol, s["pairs"])
        r = ol.oracle_greedy(b, p, threads=4, want_text=True)
        assert r["cost"].tolist() == s["cost"], s["name"]
        assert r["text"] == s["cigar"], s["name"]
        total += b.n
    assert total > 4500


def test_oracle_correction_vs_reference_fresh(ol):
    if not ol.ref_corr_available():
        pytest.skip("oracle/_ref/libgasma_ref_corr.so not built (needs /root/reference)")
    rng = np.random.default_rng(4444)
    differs = 0
    for (k, x, o, e, md) in [(3, 1, 1, 1, 12), (6, 2, 3, 1, 30), (0, 1, 1, 1, 5)]:
        b = ol.random_batch(rng, 4000, 1, 128, 0.3, 0.35, md)
        r = ol.oracle_greedy(b, ol.make_params(k, x, o, e, correction=1), threads=4, want_text=True)
        f = ol.ref_greedy(b, k, x, o, e, threads=4, correction=True)
        assert r["cost"].tolist() == f["cost"].tolist()
        assert r["text"] == f["text"]
        plain = ol.ref_greedy(b, k, x, o, e, threads=4)
        differs += sum(a != c for a, c in zip(plain["text"], f["text"]))
    assert differs > 100   # the macro really changes results (pairs with n != m)


# ---- the fast CPU arm (oracle/gasma_fast.cpp): bench.py's CPU baseline beyond the reference's 128 characters ----
def test_fast_cpu_arm_matches_port_all_widths(ol):
    rng = np.random.default_rng(777)
    for W in (1, 2, 3, 4):
        for (k, x, o, e, g) in [(3, 1, 1, 1, 1), (3, 1, 1, 1, 0), (9, 2, 3, 1, 1), (25, 1, 2, 1, 1), (50, 4, 6, 2, 1)]:
            b = ol.random_batch(rng, 1500, 1, 64 * W, 0.3, 0.3, 50)
            r = ol.oracle_greedy(b, ol.make_params(k, x, o, e, W, g), threads=4)
            f = ol.fast_greedy(b, k, x, o, e, W, g, threads=3)
            assert np.array_equal(r["cost"], f["cost"]), (W, k)
            assert np.array_equal(r["op_off"], f["op_off"]), (W, k)
            assert np.array_equal(r["ops"], f["ops"]), (W, k)
            assert r["steps"] == f["steps"]
    # non-ACGT bytes pack as 'A' (bit_convert.cpp:248), rejected pairs are flagged, empty strings work
    b = ol.Batch.from_pairs([(b"ACGTNNacgtACGT", b"ACGTAAAAAAACGT"), (b"", b""), (b"A" * 130, b"A" * 130), (b"ACGT", b"AGT")])
    r = ol.oracle_greedy(ol.Batch.from_pairs([b.pair(0), b.pair(1), b.pair(3)]), ol.make_params())
    f = ol.fast_greedy(b, W=2)
    assert f["cost"].tolist() == [r["cost"][0], r["cost"][1], -2**31, r["cost"][2]]
    assert ol.render_all(f["ops"], f["op_off"]) == ol.render_all(r["ops"], r["op_off"])[:2] + ["", "1M1I2M"]


def test_fast_cpu_arm_matches_reference_and_golden(ol, golden):
    for s in golden("greedy_ref.json")["sets"]:
        ps = s["params"]
        b = _batch(ol, s["pairs"])
        f = ol.fast_greedy(b, ps["k"], ps["x"], ps["o"], ps["e"], 2, 1, tuple(ps["probs"]), threads=2)
        assert f["cost"].tolist() == s["cost"], s["name"]
        assert ol.render_all(f["ops"], f["op_off"]) == s["cigar"], s["name"]
    if not ol.ref_available():
        pytest.skip("oracle/_ref not built (needs /root/reference)")
    rng = np.random.default_rng(778)
    b = ol.random_batch(rng, 6000, 1, 128, 0.3, 0.3, 50)
    f = ol.fast_greedy(b, 5, 2, 3, 1, 2, threads=4)
    r = ol.ref_greedy(b, 5, 2, 3, 1, threads=4)
    assert f["cost"].tolist() == r["cost"].tolist()
    assert ol.render_all(f["ops"], f["op_off"]) == r["text"]


# ---- coverage metric: the port against the unmodified benchmark_coverage.h on fresh random inputs ----
def test_coverage_port_vs_reference_fuzz(ol):
    """go_lcm / go_covers (oracle/gasma_oracle.c) against long_consecutive_matching_substring / covers
    (GASMA/benchmark/benchmark_coverage.h:26-91, unmodified, through oracle/_ref) on 100 k random CIGAR / string pairs."""
    if not ol.ref_available():
        pytest.skip("oracle/_ref not built (needs /root/reference)")
    import ctypes as C
    rng = np.random.default_rng(31337)
    letters = "MID=X"
    L, R = ol.oracle(), ol.ref()
    out_o = np.zeros(512, np.uint8)
    out_r = C.create_string_buffer(512)
    n_lcm = n_cov = 0
    for it in range(50000):
        m = int(rng.integers(0, 200))
        s1 = bytes(rng.choice(np.frombuffer(b"ACGT", np.uint8), m).tolist())
        # a CIGAR that consumes exactly the m read characters; runs of every code, lengths 1..12
        ops, left = [], m
        while left > 0 or rng.random() < 0.15:
            code = int(rng.integers(0, 5))
            ln = int(rng.integers(1, 13))
            if code != 2:
                if left == 0:
                    break
                ln = min(ln, left)
                left -= ln
            ops.append((ln << 3) | code)
        ops_np = np.array(ops if ops else [0], np.uint16)[:len(ops)]
        text = "".join("%d%s" % (o >> 3, letters[o & 7]) for o in ops).encode()
        thr = int(rng.integers(1, 5))
        n1 = L.go_lcm(ol._ptr(np.frombuffer(s1 + b"\0", np.uint8)), m, ol._ptr(ops_np) if len(ops) else None, len(ops), thr,
                      ol._ptr(out_o), 512)
        n2 = R.gasma_ref_lcm(s1, b"", text, thr, out_r, 512)
        assert n1 == n2 and out_o[:n1].tobytes() == out_r.raw[:n2], (s1, text, thr)
        n_lcm += 1
        # covers(a, b): b a subsequence of a; draw b as a noisy subsequence so that both outcomes occur
        a = out_o[:n1].tobytes()
        keep = rng.random(len(a)) < 0.6
        bb = bytearray(np.frombuffer(a, np.uint8)[keep].tolist())
        if len(bb) and rng.random() < 0.5:
            bb[int(rng.integers(0, len(bb)))] = b"ACGT"[int(rng.integers(0, 4))]
        if rng.random() < 0.1:
            bb.append(b"ACGT"[int(rng.integers(0, 4))])
        bb = bytes(bb)
        c1 = L.go_covers(ol._ptr(np.frombuffer(a + b"\0", np.uint8)), len(a), ol._ptr(np.frombuffer(bb + b"\0", np.uint8)), len(bb))
        c2 = R.gasma_ref_covers(a, bb)
        assert c1 == c2, (a, bb)
        n_cov += 1
    assert n_lcm + n_cov >= 100000


# ---- the aligner's other mask variants: flip_short_hurdles(2), flip_short_matches(1) ----
def test_oracle_mask_variants_match_golden(ol, golden):
    total = 0
    for s in golden("greedy_ref_mask.json")["sets"]:
        ps = s["params"]
        p = ol.make_params(ps["k"], ps["x"], ps["o"], ps["e"], 2, 1, tuple(ps["probs"]), mask_variant=s["variant"])
        r = ol.oracle_greedy(_batch(ol, s["pairs"]), p, threads=4, want_text=True)
        assert r["cost"].tolist() == s["cost"], (s["name"], s["variant"])
        assert r["text"] == s["cigar"], (s["name"], s["variant"])
        total += len(s["cost"])
    assert total > 4000


def test_oracle_mask_variants_vs_reference_fresh(ol):
    """Fresh fuzz against the unmodified reference whose call site hurdle_matrix.h:453 runs the other variant."""
    rng = np.random.default_rng(515)
    for variant in (1, 2):
        if not ol.ref_mask_available(variant):
            pytest.skip("oracle/_ref/libgasma_ref_mask*.so not built (needs /root/reference)")
        b = ol.random_batch(rng, 5000, 1, 128, 0.3, 0.3, 50)
        r = ol.oracle_greedy(b, ol.make_params(4, 1, 2, 1, mask_variant=variant), threads=4, want_text=True)
        f = ol.ref_greedy(b, 4, 1, 2, 1, threads=4, mask_variant=variant)
        assert r["cost"].tolist() == f["cost"].tolist()
        assert r["text"] == f["text"]
        base = ol.oracle_greedy(b, ol.make_params(4, 1, 2, 1), threads=4)
        assert int((base["cost"] != r["cost"]).sum()) > 100      # the variants really are different aligners


# ---- the synthetic workload generator (checker side; the device generator is compared with it in the GPU tests) ----
def test_generator_streams_are_independent(ol):
    """Neighbouring pairs must not replay each other's random draws (the first seeding did: pair p+1 was pair p one
    draw later, so its read was pair p's read shifted by one character)."""
    cfg = ol.gen_config(0x47415341 + 1, 150, 0.10)
    b = ol.generate(cfg, 1_000_000, 64)
    rows = [bytes(b.pair(i)[0]) for i in range(64)]
    assert all(len(r) == 150 for r in rows)
    for a, b in zip(rows, rows[1:]):
        for shift in range(0, 9):
            same = sum(1 for x, y in zip(a[shift:], b) if x == y) / (150 - shift)
            same2 = sum(1 for x, y in zip(b[shift:], a) if x == y) / (150 - shift)
            assert same < 0.5 and same2 < 0.5          # independent uniform strings agree on a quarter of the columns
    assert len(set(rows)) == 64
