"""Generates tests/golden/greedy_ref_mask.json by running the UNMODIFIED reference with the highway search on the
aligner's other mask variants (oracle/_ref/libgasma_ref_mask{1,2}.so, built by oracle/Makefile from /root/reference):
the call site hurdle_matrix.h:453 reads flip_short_hurdles(2) (variant 1) or flip_short_hurdles(1).flip_short_matches(1)
(variant 2) through a macro in oracle/ref_driver.cpp; utils.h:200-240 itself is untouched.  Canonical mode.  Run in
the build container:

    python tests/golden/make_golden_mask.py

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

LITERALS = [G1, G2, G3, ("ACGT", "ACGT"), ("ACGT", "AGT"), ("AAAA", "TTTT"), ("", ""), ("A" * 128, "A" * 128),
            (BASE102, BASE102[:20] + "GGGGG" + BASE102[20:]), ("A", ""), ("", "A"), ("ACGTACGTAC", "TTTTTTTTTTTTTTTT"),
            ("ACGTNNACGTacgtACGT", "ACGTNNACGT"), ("ACACACACACACACAC", "CACACACACACACACA"), ("T" * 128, "A" * 128)]
PARAM_SETS = [
    dict(k=3, x=1, o=1, e=1, probs=list(ol.DEFAULT_PROBS)),
    dict(k=6, x=2, o=3, e=1, probs=list(ol.DEFAULT_PROBS)),
]


def main():
    out = {"_doc": "outputs of the unmodified reference with hurdle_matrix.h:453 on another mask variant; see make_golden_mask.py",
           "sets": []}
    lit = ol.Batch.from_pairs([(a.encode(), b.encode()) for a, b in LITERALS])
    rng = np.random.default_rng(20241105)
    batches = [("literals", lit), ("gen_e10", ol.generate(ol.gen_config(0x47415361, 100, 0.10), 0, 300)),
               ("gen_e20", ol.generate(ol.gen_config(0x47415362, 100, 0.20), 0, 300)),
               ("ragged", ol.random_batch(rng, 300, 1, 128, 0.3, 0.3, 50)), ("noisy", ol.random_batch(rng, 300, 60, 128, 0.5, 0.2, 30))]
    for variant in (1, 2):
        for name, b in batches:
            for ps in PARAM_SETS:
                r = ol.ref_greedy(b, ps["k"], ps["x"], ps["o"], ps["e"], tuple(ps["probs"]), mode=0, threads=8, mask_variant=variant)
                out["sets"].append({"name": name, "variant": variant, "params": ps, "pairs": pairs_of(b),
                                    "cost": [int(c) for c in r["cost"]], "cigar": r["text"]})
    path = os.path.join(HERE, "greedy_ref_mask.json")
    with open(path, "w") as f:
        json.dump(out, f, separators=(",", ":"))
    print(path, sum(len(s["cost"]) for s in out["sets"]), "results")


if __name__ == "__main__":
    main()
