"""The bench line's contract, checked on the lines the last profile / scaling rounds committed under profiles/ (no GPU
needed): every key the driver reads is there and the numbers are consistent with each other."""
import json
import os

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _line(name):
    path = os.path.join(ROOT, "profiles", name)
    if not os.path.exists(path):
        pytest.skip("%s not committed" % name)
    with open(path) as f:
        rows = [l for l in f if l.startswith("{")]
    assert len(rows) == 1, "ONE JSON line"
    return json.loads(rows[0])


@pytest.mark.parametrize("name,n", [("r02p_bN1.json", 1), ("r02q_bN8.json", 8)])
def test_bench_line_contract(name, n):
    d = _line(name)
    for k in ("metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling", "vs_baseline",
              "dtype", "data", "config", "e2e", "gpu_launches", "clocks", "roofline", "parity"):
        assert k in d, k
    assert d["n_gpus"] == n and d["higher_is_better"] is True and d["scaling"] == "weak" and d["vs_baseline"] is None
    assert d["data"] == "synthetic" and "workload" in d["config"] and "model" not in d["config"]
    # value = pairs of all ranks / time of the K steps
    pairs = 1_000_000
    assert abs(d["value"] - n * pairs / (d["ms_per_step"] * 1e-3)) / d["value"] < 1e-6
    assert d["gpu_launches"] >= 2 * d["steps"]
    e = d["e2e"]
    assert e["h2d_bytes_per_step"] > 300e6 and e["d2h_bytes_per_step"] > 8e6 and 0 < e["value"] < d["value"]
    r = d["roofline"]
    assert r["bound"] == "hbm" and r["unit"] == "GB/s" and abs(r["frac"] - r["achieved"] / r["peak"]) < 1e-9
    assert r["kernel_ms"] < d["ms_per_step"] and abs(r["kernel_share_of_step"] - r["kernel_ms"] / d["ms_per_step"]) < 0.02
    assert abs(r["achieved"] - r["algorithmic_bytes_per_launch"] / (r["kernel_ms"] * 1e-3) / 1e9) < 1e-6 * r["achieved"]
    c = d["clocks"]
    assert c["sm_mhz"] > 0.9 * c["sm_max_mhz"] and not set(c["reasons"]) & {"hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown"}
    p = d["parity"]
    assert p["ok"] is True and p["cost_mismatch"] == 0 and p["ops_mismatch"] == 0 and p["pairs"] >= 2 * pairs
    # the other ways of driving the same calls are reported beside the plain ones and checked against them
    assert d["in_flight"]["outputs_equal_sequential"] is True and d["in_flight"]["value"] > d["value"]
    assert e["async_pair"]["equals_sync_call"] is True and d["e2e_packed"]["equals_ascii_call"] is True
    f = d["full_path"]
    assert f["parity"]["ok"] is True and f["in_flight"]["counters_equal_sequential"] is True
    assert f["counters"]["total"] == f["counters"]["nw_correct"] == n * pairs * f["steps"]
    for leg in d["configs"].values():
        assert leg["parity"]["ok"] is True and leg["roofline"]["kernel_ms"] <= leg["ms_per_step"]
    if n == 1:
        b = d["cpu_baseline"]
        assert b["kind"] in ("reference", "port") and b["cores"] >= 1 and b["value"] > 0 and "sample" in b
        assert r["traffic"] is None or r["traffic"] > r["algorithmic_bytes_per_launch"]


def test_traffic_json_belongs_to_these_sources():
    """profiles/traffic.json is keyed by the hash of the sources the profiled library was built from: after any change to
    the CUDA sources the profile round has to be repeated (tools/profile_round.sh), or bench.py drops the figures."""
    import importlib.util
    spec = importlib.util.spec_from_file_location("gasma_b200_build", os.path.join(ROOT, "approximate-string-matching_b200", "build.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
        t = json.load(f)
    assert t["source_sha1"] == mod.source_sha()
    assert set(t["legs"]) >= {"headline", "full_path"}
    assert t["legs"]["headline"]["traffic_bytes"] > 3e8 and t["legs"]["headline"]["warp_instructions"] > 1e8
