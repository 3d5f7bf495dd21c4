"""Generates tests/golden/*.json by running the UNMODIFIED reference (oracle/_ref/libgasma_ref.so, built by
oracle/Makefile from /root/reference) in canonical mode.  Run in the build container:

    python tests/golden/make_golden.py

The fixtures are committed; /root/reference does not exist on the GPU box, so tests only read the JSON.
Inputs of the bulk sets come from the seeded generator (oracle twin go_generate_pair) and from numpy's
PCG64 with fixed seeds; both are stored verbatim so the fixtures do not depend on either staying stable.
"""
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import oracle_lib as ol  # noqa: E402

# literal pairs taken from the reference tree
G1 = ("TGCGTCGCAGTGTCGAAAAAAATCTAAAGGGCCTACCGGGCCCCTTCATGAAGAGCAGGGAAAGAATCTCTATACTATTACTAATGGAAAGCAGTCTTGG",
      "TGCGTCGCAGTGTCGAAATAAAATCTAAAGGGCCTACCGGGCCCCTTCATGAAGAGTAGGTAAAGAATCTCTAGTACTACTACTAATGGAAAGCAGTCTTGG")  # GASMA/main.cpp:7-8
G2 = ("GCCCCGTCCCAGCACAGGCAGCGGGGATGCTCTAGAGCGATGTCGACCTGGGAAAGGCGCTGGGGCGCGCCGAATTCCAAAGGAGTTCCGTAAGGTTCAG",
      "GCTCCTCCCAGCACAGGCAGCGGGATGCTCTAGAGCATTCGCCCTGGGAAGGCGTGGGGCCCCCTAATTCCAAAAGGAGTTGCCCGTAAGGTTCAG")  # GASMA/main.cpp:14-15
G3 = ("TAAAGGAATAACCCCGCACAGCGTCGTGGCTGGTGAGTCGTCTATATACAATGTGTCCCATCGCACGTATCTCAAACTAGACGGCACTGTTCAGTCTCAA",
      "TAAAGAATAACCCCGCACAGCGTCGTGGCTGGTGAGTTGTCTATATACATTTGTGTCCCATCGCACGTATCTCAAACTAGACGTGCACTGTTCAGTCTCAA")  # GASMA/benchmark/notes:1-2, dashes removed
BASE102 = (G1[1])[:102]
LITERALS = [G1, G2, G3, ("ACGT", "ACGT"), ("ACGT", "AGT"), ("AAAA", "TTTT"), ("", ""), ("A" * 128, "A" * 128),
            (BASE102, BASE102[:20] + "GGGGG" + BASE102[20:]),  # destination lane outside the band (n-m=5 > k=3)
            ("ACGTNNACGTacgtACGT", "ACGTNNACGTacgtACGT"), ("NNNNNNNN", "AAAAAAAA"), ("A", ""), ("", "A"),
            ("ACGTACGTAC", "TTTTTTTTTTTTTTTT"), ("C" * 128, "C" * 78)]

PARAM_SETS = [
    dict(k=3, x=1, o=1, e=1, probs=list(ol.DEFAULT_PROBS)),          # benchmark.cpp:22
    dict(k=5, x=1, o=1, e=1, probs=list(ol.DEFAULT_PROBS)),
    dict(k=3, x=2, o=3, e=1, probs=list(ol.DEFAULT_PROBS)),
    dict(k=8, x=4, o=6, e=2, probs=[0.95, 0.02, 0.03]),             # hurdle_matrix.h:481-483 defaults
    dict(k=50, x=1, o=2, e=1, probs=list(ol.DEFAULT_PROBS)),
]


def run_ref(batch, ps):
    r = ol.ref_greedy(batch, ps["k"], ps["x"], ps["o"], ps["e"], tuple(ps["probs"]), mode=0, threads=8)
    return [int(c) for c in r["cost"]], r["text"]


def pairs_of(batch):
    return [[a.decode("latin1"), b.decode("latin1")] for a, b in (batch.pair(i) for i in range(batch.n))]


def main():
    out = {"_doc": "outputs of the unmodified reference in canonical mode; see make_golden.py", "sets": []}
    lit = ol.Batch.from_pairs([(a.encode(), b.encode()) for a, b in LITERALS])
    for ps in PARAM_SETS:
        cost, text = run_ref(lit, ps)
        out["sets"].append({"name": "literals", "params": ps, "pairs": pairs_of(lit), "cost": cost, "cigar": text})
    # seeded generator sets: the BASELINE configs at small n (L=100; e = 0.05, 0.10, 0.20 exact; Illumina profile)
    gens = [("gen_e05", ol.gen_config(0x47415341, 100, 0.05)), ("gen_e10", ol.gen_config(0x47415342, 100, 0.10)),
            ("gen_e20", ol.gen_config(0x47415343, 100, 0.20)),
            ("gen_e20_lt", ol.gen_config(0x47415344, 100, 0.20, mode=1)),
            ("gen_illumina", ol.gen_config(0x47415345, 100, 0.0, mode=2, profile=(0.02452, 0.000468, 0.000553))),
            ("gen_L128_e10", ol.gen_config(0x47415346, 118, 0.10))]
    for name, cfg in gens:
        b = ol.generate(cfg, 0, 400)
        for ps in PARAM_SETS[:2]:
            cost, text = run_ref(b, ps)
            out["sets"].append({"name": name, "params": ps, "pairs": pairs_of(b), "cost": cost, "cigar": text})
    # ragged random sets (lengths 1..128, up to 30 % errors, many indels, |n-m| up to 50)
    rng = np.random.default_rng(20240607)
    for name, args in [("ragged", (1, 128, 0.3, 0.3, 50)), ("short", (1, 12, 0.5, 0.5, 12)), ("indel_heavy", (90, 128, 0.2, 0.6, 50))]:
        b = ol.random_batch(rng, 300, *args)
        for ps in PARAM_SETS:
            cost, text = run_ref(b, ps)
            out["sets"].append({"name": name, "params": ps, "pairs": pairs_of(b), "cost": cost, "cigar": text})
    path = os.path.join(HERE, "greedy_ref.json")
    with open(path, "w") as f:
        json.dump(out, f, separators=(",", ":"))
    print("wrote", path, os.path.getsize(path), "bytes,", sum(len(s["cost"]) for s in out["sets"]), "results")

    # sse3_convert2bit1 fixtures: plane bytes + the in-place permutation
    conv = []
    rng = np.random.default_rng(99)
    alpha = np.frombuffer(b"ACGTNacgt\0", dtype=np.uint8)
    for i in range(24):
        s = alpha[rng.integers(0, 4 if i < 12 else len(alpha), 128)].tobytes()
        b0, b1, sh = ol.ref_convert2bit1(s)
        conv.append({"in": s.decode("latin1"), "bits0": b0.tolist(), "bits1": b1.tolist(), "shuffled": sh.tobytes().decode("latin1")})
    sig = np.zeros(3)
    ol.ref().gasma_ref_significance(*ol.DEFAULT_PROBS, ol._ptr(sig))
    with open(os.path.join(HERE, "convert2bit_ref.json"), "w") as f:
        json.dump({"vectors": conv, "significance_default": [float.hex(float(v)) for v in sig]}, f, separators=(",", ":"))

    # coverage metric fixtures from the unmodified benchmark_coverage.h
    cov = []
    for (a, b), cig in [(G1, "22M1D50M1D28M"), (G3, "7M1I44M1D31M1D17M"), (("ACGTACGT", "ACGTACGT"), "3=1X4="),
                        (("ACGTACGT", "ACTACGT"), "2=1I5="), (("ACGT", "ACGT"), "2M2M")]:
        import ctypes as C
        for thr in (1, 3):
            buf = C.create_string_buffer(600)
            ol.ref().gasma_ref_lcm(a.encode(), b.encode(), cig.encode(), thr, buf, 600)
            cov.append({"s1": a, "s2": b, "cigar": cig, "threshold": thr, "lcm": buf.value.decode()})
    cv = [{"a": a, "b": b, "covers": int(ol.ref().gasma_ref_covers(a.encode(), b.encode()))}
          for a, b in [("ACGT", "AGT"), ("ACGT", "ACGT"), ("AGT", "ACGT"), ("", ""), ("ACGT", ""), ("AC", "CA"), ("AACC", "AC")]]
    with open(os.path.join(HERE, "coverage_ref.json"), "w") as f:
        json.dump({"lcm": cov, "covers": cv}, f, separators=(",", ":"))


if __name__ == "__main__":
    main()
