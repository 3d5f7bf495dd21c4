"""CPU tests of the kernels' per-pair __host__ __device__ code (compiled with g++ by the host_sim fixture)
against the oracle and the golden vectors.  This is how the device logic is checked in the GPU-less build
container; the -m gpu tests repeat the comparison through the real C ABI on a B200."""
import numpy as np
import pytest

import sim_lib


def _batch(ol, pairs):
    return ol.Batch.from_pairs([(a.encode("latin1"), b.encode("latin1")) for a, b in pairs])


def test_kernel_logic_matches_golden(ol, host_sim, golden):
    for s in golden("greedy_ref.json")["sets"]:
        ps = s["params"]
        p = ol.make_params(ps["k"], ps["x"], ps["o"], ps["e"], 2, 1, tuple(ps["probs"]))
        b = _batch(ol, s["pairs"])
        cost, text, st, _ = sim_lib.sim_greedy(host_sim, ol, b, p, 128 if ps["k"] < 18 else 32)
        assert st == 0
        assert cost.tolist() == s["cost"], s["name"]
        assert text == s["cigar"], s["name"]


@pytest.mark.parametrize("W", [1, 2, 3, 4])
def test_kernel_logic_matches_oracle_all_widths(ol, host_sim, W):
    rng = np.random.default_rng(100 + W)
    for (k, x, o, e, g) in [(3, 1, 1, 1, 1), (7, 2, 3, 1, 1), (3, 1, 1, 1, 0), (0, 1, 1, 1, 1)]:
        b = ol.random_batch(rng, 1500, 1, 64 * W, 0.3, 0.3, 50)
        p = ol.make_params(k, x, o, e, W, g)
        r = ol.oracle_greedy(b, p, threads=4, want_text=True)
        cost, text, st, steps = sim_lib.sim_greedy(host_sim, ol, b, p)
        assert st == 0
        assert cost.tolist() == r["cost"].tolist()
        assert text == r["text"]
        assert steps == r["steps"]


def test_pack_logic_matches_golden_and_oracle(ol, host_sim, golden):
    vec = golden("convert2bit_ref.json")["vectors"]
    seqs = np.frombuffer("".join(v["in"] for v in vec).encode("latin1"), np.uint8)
    off = (np.arange(len(vec) + 1) * 128).astype(np.uint32)
    p0, p1, st = sim_lib.sim_pack(host_sim, seqs, off, 2)
    assert st == 0
    for i, v in enumerate(vec):
        assert p0[i].view(np.uint8).tolist() == v["bits0"]
        assert p1[i].view(np.uint8).tolist() == v["bits1"]
    # ragged lengths, arbitrary bytes, every alignment of string starts within 16-byte chunks
    rng = np.random.default_rng(5)
    for W in (1, 2, 3, 4):
        lens = rng.integers(0, 64 * W + 1, 700)
        off = np.zeros(len(lens) + 1, np.uint32)
        off[1:] = np.cumsum(lens)
        seqs = rng.choice(np.frombuffer(b"ACGTNacgt\0\xff", np.uint8), int(off[-1])).astype(np.uint8)
        p0, p1, st = sim_lib.sim_pack(host_sim, seqs, off, W, 128)
        assert st == 0
        for i in range(len(lens)):
            e0, e1 = ol.oracle_pack(seqs[off[i]:off[i + 1]].tobytes(), W)
            assert p0[i].tolist() == e0.tolist() and p1[i].tolist() == e1.tolist()


def test_length_and_band_violations_are_flagged(ol, host_sim):
    b = _batch(ol, [("A" * 129, "A" * 100), ("ACGT", "ACGT")])
    cost, _, st, _ = sim_lib.sim_greedy(host_sim, ol, b, ol.make_params())
    assert st & 1 and cost[0] == -2**31 and cost[1] == 0
    b = _batch(ol, [("A" * 10, "A" * 61)])
    _, _, st, _ = sim_lib.sim_greedy(host_sim, ol, b, ol.make_params())
    assert st & 2


@pytest.mark.parametrize("W", [2, 3])
def test_widened_band_logic_matches_oracle(ol, host_sim, W):
    """KParams::correction (the reference's -DCORRECTION band, hurdle_matrix.h:634-641) vs the oracle in that mode."""
    rng = np.random.default_rng(300 + W)
    for (k, x, o, e, g, md) in [(3, 1, 1, 1, 1, 12), (5, 2, 3, 1, 1, 20), (3, 1, 1, 1, 0, 9), (0, 1, 1, 1, 1, 6)]:
        b = ol.random_batch(rng, 1500, 1, 64 * W, 0.3, 0.35, md)
        p = ol.make_params(k, x, o, e, W, g, correction=1)
        r = ol.oracle_greedy(b, p, threads=4, want_text=True)
        cost, text, st, steps = sim_lib.sim_greedy(host_sim, ol, b, p, max_delta=md)
        assert st == 0
        assert cost.tolist() == r["cost"].tolist()
        assert text == r["text"]
        assert steps == r["steps"]
    # a pair beyond the lane slots is flagged, not silently narrowed
    b = _batch(ol, [("A" * 10, "A" * 20), ("ACGT", "ACGT")])
    cost, _, st, _ = sim_lib.sim_greedy(host_sim, ol, b, ol.make_params(correction=1), max_delta=8)
    assert st & 2 and cost[0] == -2**31 and cost[1] == 0


def test_widened_band_logic_matches_golden(ol, host_sim, golden):
    for s in golden("greedy_ref_correction.json")["sets"]:
        ps = s["params"]
        p = ol.make_params(ps["k"], ps["x"], ps["o"], ps["e"], 2, 1, tuple(ps["probs"]), correction=1)
        b = _batch(ol, s["pairs"])
        cost, text, st, _ = sim_lib.sim_greedy(host_sim, ol, b, p, max_delta=ps["max_delta"])
        assert st == 0
        assert cost.tolist() == s["cost"], s["name"]
        assert text == s["cigar"], s["name"]


def test_numa_binding_helper(pkg, tmp_path, monkeypatch):
    """shard.bind_host_to_device on a fake sysfs tree: picks the GPU's node, intersects with the current affinity,
    and leaves single-node boxes alone."""
    import os
    import types
    from gasma_b200 import shard
    import torch
    props = types.SimpleNamespace(pci_domain_id=0, pci_bus_id=0x1b, pci_device_id=0)
    monkeypatch.setattr(torch.cuda, "get_device_properties", lambda d: props)
    sysfs = tmp_path / "sys"
    (sysfs / "bus/pci/devices/0000:1b:00.0").mkdir(parents=True)
    (sysfs / "bus/pci/devices/0000:1b:00.0/numa_node").write_text("1\n")
    for node, cpus in ((0, "0-1"), (1, "%s" % ",".join(str(c) for c in sorted(os.sched_getaffinity(0))[:1]))):
        (sysfs / ("devices/system/node/node%d" % node)).mkdir(parents=True)
        (sysfs / ("devices/system/node/node%d/cpulist" % node)).write_text(cpus + "\n")
    assert shard.device_numa_node(0, str(sysfs)) == 1
    before = os.sched_getaffinity(0)
    try:
        got = shard.bind_host_to_device(0, str(sysfs))
        assert got == (1, 1) and len(os.sched_getaffinity(0)) == 1
    finally:
        os.sched_setaffinity(0, before)
    # one node only: nothing to bind to
    import shutil
    shutil.rmtree(sysfs / "devices/system/node/node0")
    assert shard.bind_host_to_device(0, str(sysfs)) is None
    assert shard._parse_cpulist("0-3,8,10-11") == {0, 1, 2, 3, 8, 10, 11}


@pytest.mark.parametrize("variant", [1, 2])
def test_mask_variant_logic_matches_oracle_and_golden(ol, host_sim, golden, variant):
    """flip_word<FLIP> (greedy.cuh) -- flip_short_hurdles(2) and flip_short_hurdles(1).flip_short_matches(1), rebuilt word
    by word inside the highway scans -- against the oracle for W = 1..4 and against the unmodified reference's outputs."""
    rng = np.random.default_rng(640 + variant)
    for W in (1, 2, 3, 4):
        for (k, x, o, e, g) in [(3, 1, 1, 1, 1), (7, 2, 3, 1, 1), (3, 1, 1, 1, 0)]:
            b = ol.random_batch(rng, 1200, 1, 64 * W, 0.35, 0.3, 50)
            p = ol.make_params(k, x, o, e, W, g, mask_variant=variant)
            r = ol.oracle_greedy(b, p, threads=4, want_text=True)
            cost, text, st, steps = sim_lib.sim_greedy(host_sim, ol, b, p, mask_variant=variant)
            assert st == 0
            assert cost.tolist() == r["cost"].tolist(), (W, k)
            assert text == r["text"], (W, k)
            assert steps == r["steps"]
    for s in golden("greedy_ref_mask.json")["sets"]:
        if s["variant"] != variant:
            continue
        ps = s["params"]
        p = ol.make_params(ps["k"], ps["x"], ps["o"], ps["e"], 2, 1, tuple(ps["probs"]), mask_variant=variant)
        cost, text, st, _ = sim_lib.sim_greedy(host_sim, ol, _batch(ol, s["pairs"]), p, mask_variant=variant)
        assert st == 0 and cost.tolist() == s["cost"] and text == s["cigar"], s["name"]
