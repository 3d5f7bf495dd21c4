"""Generates tests/golden/leap_ref.json: penalties of the UNMODIFIED LEAP comparator (class LV of
GASMA/benchmark/LEAP_SIMD/LV_BAG.cpp, driven like benchmark::_run_LEAP; compiled into oracle/_ref by
oracle/Makefile) for the input pairs already stored in greedy_ref.json.  Run in the build container:

    python tests/golden/make_golden_leap.py

-2147483648 marks a pair for which no lane converges within af_threshold = 200 (the reference then reports the
previous pair's value)."""
import json
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import oracle_lib as ol  # noqa: E402

PARAMS = [(3, 1, 1, 1), (5, 1, 1, 1), (3, 2, 3, 1), (8, 4, 6, 2)]   # (k, x, o, e); benchmark.cpp:22 uses the first


def main():
    with open(os.path.join(HERE, "greedy_ref.json")) as f:
        g = json.load(f)
    out = {"_doc": "see make_golden_leap.py; set_index refers to greedy_ref.json['sets']", "sets": []}
    seen = set()
    for si, s in enumerate(g["sets"]):
        key = (s["name"], len(s["pairs"]))
        if key in seen:      # the same pairs appear once per greedy parameter set
            continue
        seen.add(key)
        b = ol.Batch.from_pairs([(a.encode("latin1"), c.encode("latin1")) for a, c in s["pairs"]])
        for (k, x, o, e) in PARAMS:
            pen, _ = ol.ref_leap(b, k, x, o, e, mode=0, threads=1)
            out["sets"].append({"set_index": si, "name": s["name"], "k": k, "x": x, "o": o, "e": e, "penalty": [int(v) for v in pen]})
    path = os.path.join(HERE, "leap_ref.json")
    with open(path, "w") as f:
        json.dump(out, f, separators=(",", ":"))
    print("wrote", path, os.path.getsize(path), "bytes,", sum(len(s["penalty"]) for s in out["sets"]), "results")


if __name__ == "__main__":
    main()
