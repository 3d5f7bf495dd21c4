"""GPU parity tests for the LEAP comparator kernel through the C ABI."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _batch(ol, pairs):
    return ol.Batch.from_pairs([(a.encode("latin1"), b.encode("latin1")) for a, b in pairs])


def test_gpu_leap_matches_golden(pkg, ol, golden):
    """Penalties captured from the unmodified class LV (tests/golden/make_golden_leap.py)."""
    g = golden("greedy_ref.json")
    lp = golden("leap_ref.json")
    aligners = {}
    total = 0
    for s in lp["sets"]:
        pairs = g["sets"][s["set_index"]]["pairs"]
        if any(not set(a + c) <= set("ACGT") for a, c in pairs):
            continue      # the 2-bit path maps other characters to 'A'; LV compares raw bytes (documented domain: ACGT)
        key = (s["k"], s["x"], s["o"], s["e"])
        if key not in aligners:
            aligners[key] = pkg.Aligner(k=s["k"], x=s["x"], o=s["o"], e=s["e"], max_length=128)
        b = _batch(ol, pairs)
        pen = aligners[key].leap_batch(b.reads, b.read_off, b.refs, b.ref_off)
        assert pen.tolist() == s["penalty"], (s["name"], key)
        total += b.n
    assert total > 9000
    for al in aligners.values():
        al.close()


@pytest.mark.parametrize("W", [1, 2, 3, 4])
def test_gpu_leap_matches_oracle(pkg, ol, W):
    rng = np.random.default_rng(900 + W)
    for (k, x, o, e) in [(3, 1, 1, 1), (6, 1, 2, 1), (3, 2, 3, 1), (10, 4, 6, 2)]:
        al = pkg.Aligner(k=k, x=x, o=o, e=e, max_length=64 * W)
        for max_err, indel in ((0.1, 0.2), (0.9, 0.5)):
            b = ol.random_batch(rng, 4000, 0, 64 * W, max_err, indel, 50)
            exp = ol.oracle_leap(b, k, x, o, e, threads=8)
            assert al.leap_batch(b.reads, b.read_off, b.refs, b.ref_off).tolist() == exp.tolist()
        al.close()


def test_gpu_leap_accuracy_on_benchmark_data(pkg, ol):
    """README.md:17,64: LEAP matches the optimum on 99.757 % of the pairs at e = 0.05 and 88.579 % at e = 0.20."""
    import torch
    al = pkg.Aligner()
    for e, lo, hi in ((0.05, 0.99, 1.0), (0.20, 0.84, 0.93)):
        n = 30000
        reads, roff, refs, foff = al.generate(pkg.gen_config(0x47415360, 100, e), 0, n)
        leap = al.leap_batch(reads, roff, refs, foff)
        nw = al.nw_batch(reads, roff, refs, foff, want_ops=False)["penalty"]
        b = ol.Batch(reads, roff, refs, foff)
        assert leap.tolist() == ol.oracle_leap(b, threads=8).tolist()
        acc = float((leap == nw).mean())
        assert lo <= acc <= hi, acc
        # device entry point
        d = [torch.from_numpy(np.ascontiguousarray(x)).cuda() for x in (np.concatenate([reads, np.zeros(16, np.uint8)]),
                                                                       roff.view(np.int32), np.concatenate([refs, np.zeros(16, np.uint8)]),
                                                                       foff.view(np.int32))]
        d_pen = torch.empty(n, dtype=torch.int32, device="cuda")
        al.leap_batch_dev(d[0], d[1], d[2], d[3], n, d_pen)
        al.synchronize()
        assert d_pen.cpu().numpy().tolist() == leap.tolist()
        ms, launches = al.kernel_time_ms(2)
        assert launches >= 1 and ms > 0
    al.close()


def test_gpu_leap_rejects_what_it_cannot_do(pkg):
    al = pkg.Aligner(x=1, o=1, e=2)    # extension dearer than opening: the reference indexes out of bounds
    reads = np.frombuffer(b"ACGT", np.uint8)
    off = np.array([0, 4], np.uint32)
    with pytest.raises(Exception):
        al.leap_batch(reads, off, reads, off)
    al.close()
