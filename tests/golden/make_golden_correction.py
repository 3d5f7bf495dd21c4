"""Generates tests/golden/greedy_ref_correction.json by running the UNMODIFIED reference compiled with its own
compile-time switch -DCORRECTION (oracle/_ref/libgasma_ref_corr.so, built by oracle/Makefile from /root/reference;
hurdle_matrix.h:501-512, :634-641) in canonical mode.  Run in the build container:

    python tests/golden/make_golden_correction.py

The fixture is committed; /root/reference does not exist on the GPU box, so tests only read the JSON.
"""
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, HERE)
import oracle_lib as ol  # noqa: E402
from make_golden import G1, G2, G3, BASE102, pairs_of  # noqa: E402

# |n-m| + k must stay <= MAX_K = 50 (the reference's lane arrays); max_delta = lane slots the GPU context needs
LITERALS = [G1, G2, G3, ("ACGT", "ACGT"), ("ACGT", "AGT"), ("AAAA", "TTTT"), ("", ""), ("A" * 128, "A" * 128),
            (BASE102, BASE102[:20] + "GGGGG" + BASE102[20:]), (BASE102[:20] + "GGGGG" + BASE102[20:], BASE102),
            ("A", ""), ("", "A"), ("ACGTACGTAC", "TTTTTTTTTTTTTTTT"), ("C" * 128, "C" * 100), ("ACGTNNACGTacgtACGT", "ACGTNNACGT")]
PARAM_SETS = [
    dict(k=3, x=1, o=1, e=1, probs=list(ol.DEFAULT_PROBS), max_delta=28),
    dict(k=5, x=1, o=1, e=1, probs=list(ol.DEFAULT_PROBS), max_delta=28),
    dict(k=3, x=2, o=3, e=1, probs=list(ol.DEFAULT_PROBS), max_delta=28),
    dict(k=8, x=4, o=6, e=2, probs=[0.95, 0.02, 0.03], max_delta=28),
]


def run_ref(batch, ps):
    r = ol.ref_greedy(batch, ps["k"], ps["x"], ps["o"], ps["e"], tuple(ps["probs"]), mode=0, threads=8, correction=True)
    return [int(c) for c in r["cost"]], r["text"]


def main():
    out = {"_doc": "outputs of the unmodified reference built with -DCORRECTION, canonical mode; see make_golden_correction.py",
           "sets": []}
    lit = ol.Batch.from_pairs([(a.encode(), b.encode()) for a, b in LITERALS])
    for ps in PARAM_SETS:
        cost, text = run_ref(lit, ps)
        out["sets"].append({"name": "literals", "params": ps, "pairs": pairs_of(lit), "cost": cost, "cigar": text})
    # generator sets with a large indel share (the widened band only matters when n != m)
    gens = [("gen_e10_indel", ol.gen_config(0x47415351, 100, 0.10, mismatch_rate=0.5)),
            ("gen_e20_indel", ol.gen_config(0x47415352, 100, 0.20, mismatch_rate=0.6)),
            ("gen_e20", ol.gen_config(0x47415353, 100, 0.20))]
    for name, cfg in gens:
        b = ol.generate(cfg, 0, 300)
        for ps in PARAM_SETS[:2]:
            cost, text = run_ref(b, ps)
            out["sets"].append({"name": name, "params": ps, "pairs": pairs_of(b), "cost": cost, "cigar": text})
    rng = np.random.default_rng(20240611)
    for name, args in [("ragged", (1, 128, 0.3, 0.3, 28)), ("short", (1, 12, 0.5, 0.5, 12)), ("indel_heavy", (90, 128, 0.2, 0.6, 28))]:
        b = ol.random_batch(rng, 250, *args)
        for ps in PARAM_SETS:
            cost, text = run_ref(b, ps)
            out["sets"].append({"name": name, "params": ps, "pairs": pairs_of(b), "cost": cost, "cigar": text})
    path = os.path.join(HERE, "greedy_ref_correction.json")
    with open(path, "w") as f:
        json.dump(out, f, separators=(",", ":"))
    print("wrote", path, os.path.getsize(path), "bytes,", sum(len(s["cost"]) for s in out["sets"]), "results")


if __name__ == "__main__":
    main()
