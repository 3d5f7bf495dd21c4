"""GPU parity tests for K4 (NW verifier) and K5 (accuracy / coverage counters) through the C ABI."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("W", [1, 2, 3, 4])
def test_gpu_nw_matches_oracle(pkg, ol, W):
    rng = np.random.default_rng(700 + W)
    for (x, o, e) in [(1, 1, 1), (2, 3, 1), (4, 6, 2)]:
        b = ol.random_batch(rng, 3000, 1, 64 * W, 0.3, 0.4, 50)
        al = pkg.Aligner(k=3, x=x, o=o, e=e, max_length=64 * W)
        r = ol.oracle_nw(b, x, o, e, threads=8)
        g = al.greedy_batch(b.reads, b.read_off, b.refs, b.ref_off, want_ops=False)["cost"]
        for bound in (None, g):           # full matrix (wide kernel) and band from the greedy penalty
            out = al.nw_batch(b.reads, b.read_off, b.refs, b.ref_off, bound=bound)
            assert out["penalty"].tolist() == r["penalty"].tolist()
            assert out["op_off"].tolist() == r["op_off"].tolist()
            assert out["ops"][:out["n_ops"]].tolist() == r["ops"].tolist()
        out = al.nw_batch(b.reads, b.read_off, b.refs, b.ref_off, bound=g, want_ops=False)
        assert out["penalty"].tolist() == r["penalty"].tolist()
        al.close()


def test_gpu_benchmark_counters(pkg, ol):
    """greedy + NW + counters on the README setup (L=100, e in {0.05, 0.20}) == oracle's counters."""
    import torch
    al = pkg.Aligner()
    for e in (0.05, 0.20):
        n = 20000
        cfg = pkg.gen_config(0x47415350, 100, e)
        d_reads = torch.empty(n * 100 + 16, dtype=torch.uint8, device="cuda")
        d_refs = torch.empty(al.generate_ref_capacity(cfg, n), dtype=torch.uint8, device="cuda")
        d_roff = torch.empty(n + 1, dtype=torch.int32, device="cuda")
        d_foff = torch.empty(n + 1, dtype=torch.int32, device="cuda")
        al.generate_dev(cfg, 0, n, d_reads, d_roff, d_refs, d_foff)
        g_cost = torch.empty(n, dtype=torch.int32, device="cuda")
        g_ops = torch.empty(n * 64, dtype=torch.int16, device="cuda")
        g_off = torch.empty(n + 1, dtype=torch.int32, device="cuda")
        al.greedy_batch_dev(d_reads, d_roff, d_refs, d_foff, n, g_cost, g_ops, g_off)
        nw_pen = torch.empty(n, dtype=torch.int32, device="cuda")
        nw_ops = torch.empty(n * 128, dtype=torch.int16, device="cuda")
        nw_off = torch.empty(n + 1, dtype=torch.int32, device="cuda")
        al.nw_batch_dev(d_reads, d_roff, d_refs, d_foff, n, g_cost, nw_pen, nw_ops, nw_off)
        counters = torch.zeros(4, dtype=torch.int64, device="cuda")
        al.metrics_dev(d_reads, d_roff, n, g_cost, g_ops, g_off, nw_pen, nw_ops, nw_off, None, counters)
        al.metrics_dev(d_reads, d_roff, n, g_cost, g_ops, g_off, nw_pen, nw_ops, nw_off, None, counters)  # accumulates
        al.synchronize()
        b = ol.generate(ol.gen_config(0x47415350, 100, e), 0, n)
        og = ol.oracle_greedy(b, ol.make_params(), threads=8)
        on = ol.oracle_nw(b, 1, 1, 1, threads=8)
        exp = ol.oracle_metrics(b, og, on)
        assert nw_pen.cpu().numpy().tolist() == on["penalty"].tolist()
        assert counters.cpu().numpy().tolist() == [2 * int(v) for v in exp]
        acc = exp[2] / exp[0]
        assert (0.85 < acc < 0.99) if e == 0.05 else (0.30 < acc < 0.70)   # README.md:18,65: 92.975 % / 46.023 %


def test_gpu_benchmark_batch_host_call(pkg, ol):
    """gasma_benchmark_batch (host pointers): the whole _run_benchmark body in one call == oracle counters,
    with and without an answers file, accumulating like the class members do."""
    al = pkg.Aligner()
    cfg = pkg.gen_config(0x47415351, 100, 0.15)
    reads, roff, refs, foff = al.generate(cfg, 1000, 30000)
    b = ol.Batch(reads, roff, refs, foff)
    tw = ol.generate(ol.gen_config(0x47415351, 100, 0.15), 1000, 30000)
    assert b.reads.tolist() == tw.reads.tolist() and b.refs.tolist() == tw.refs.tolist()   # host generate == CPU twin
    og = ol.oracle_greedy(b, ol.make_params(), threads=8)
    on = ol.oracle_nw(b, 1, 1, 1, threads=8)
    exp = [int(v) for v in ol.oracle_metrics(b, og, on)]
    out = al.benchmark_batch(b.reads, b.read_off, b.refs, b.ref_off, want_penalties=True)
    assert out["counters"].tolist() == exp
    assert out["greedy_cost"].tolist() == og["cost"].tolist()
    assert out["nw_penalty"].tolist() == on["penalty"].tolist()
    assert out["device_ms"][0] > 0 and out["device_ms"][1] > 0
    # answers: half the pairs get a wrong "optimal" penalty, the rest INT32_MIN (= use NW)
    ans = np.full(b.n, np.iinfo(np.int32).min, np.int32)
    ans[::2] = on["penalty"][::2] + 1
    exp2 = [int(v) for v in ol.oracle_metrics(b, og, on, ans)]
    out2 = al.benchmark_batch(b.reads, b.read_off, b.refs, b.ref_off, answers=ans, counters=out["counters"])
    assert out2["counters"].tolist() == [a + c for a, c in zip(exp, exp2)]
    al.close()


@pytest.mark.parametrize("W,penalties", [(1, (1, 1, 1)), (2, (1, 1, 1)), (4, (1, 1, 1)), (2, (2, 3, 1))])
def test_gpu_benchmark_batch_noisy_pairs(pkg, ol, W, penalties):
    """The fused verifier + counters kernel (unit penalties) including the pairs it defers to the Gotoh kernel
    (very noisy pairs: band > 32 diagonals), and the separate verifier / counter kernels for other penalties."""
    x, o, e = penalties
    rng = np.random.default_rng(800 + W + x)
    b = ol.random_batch(rng, 6000, 0, 64 * W, 0.6, 0.4, 50)
    og = ol.oracle_greedy(b, ol.make_params(3, x, o, e, W, 1), threads=8)
    on = ol.oracle_nw(b, x, o, e, threads=8)
    exp = [int(v) for v in ol.oracle_metrics(b, og, on)]
    al = pkg.Aligner(x=x, o=o, e=e, max_length=64 * W)
    out = al.benchmark_batch(b.reads, b.read_off, b.refs, b.ref_off, want_penalties=True)
    assert out["greedy_cost"].tolist() == og["cost"].tolist()
    assert out["nw_penalty"].tolist() == on["penalty"].tolist()
    assert out["counters"].tolist() == exp
    assert 0 < exp[3] < b.n
    al.close()


@pytest.mark.parametrize("mode", ["semi_global", "out_of_band", "general_penalties"])
def test_gpu_benchmark_batch_when_greedy_is_no_upper_bound(pkg, ol, mode):
    """The verifier sizes its band from the greedy penalty, which is NOT an upper bound on the optimum in SEMI_GLOBAL
    mode (lane switches at the ends are free) nor for pairs with |n-m| > k (the out-of-band destination lane counts
    no hurdles).  A banded result above the bound exposes that, and the pair is redone from its exact optimum: NW
    penalties, their tracebacks (through the coverage counter) and all counters must still equal the oracle's."""
    rng = np.random.default_rng({"semi_global": 911, "out_of_band": 912, "general_penalties": 913}[mode])
    x, o, e = (2, 3, 1) if mode == "general_penalties" else (1, 1, 1)
    g = 0 if mode == "semi_global" else 1
    # indel-heavy pairs with length differences well beyond k = 3
    b = ol.random_batch(rng, 20000, 20, 128, 0.25, 0.7 if mode != "semi_global" else 0.4, 40)
    og = ol.oracle_greedy(b, ol.make_params(3, x, o, e, 2, g), threads=8)
    on = ol.oracle_nw(b, x, o, e, threads=8)
    below = int((og["cost"] < on["penalty"]).sum())
    assert below > 0, "the case must contain pairs whose greedy penalty is below the optimum"
    exp = [int(v) for v in ol.oracle_metrics(b, og, on)]
    al = pkg.Aligner(x=x, o=o, e=e, max_length=128, alignment_type=1 - g)
    out = al.benchmark_batch(b.reads, b.read_off, b.refs, b.ref_off, want_penalties=True)
    assert out["greedy_cost"].tolist() == og["cost"].tolist()
    assert int((out["nw_penalty"] != on["penalty"]).sum()) == 0
    assert out["counters"].tolist() == exp
    # the same through the stand-alone verifier call with the greedy penalties as (invalid) bounds
    nw = al.nw_batch(b.reads, b.read_off, b.refs, b.ref_off, bound=og["cost"], want_ops=True)
    assert int((nw["penalty"] != on["penalty"]).sum()) == 0
    assert np.array_equal(nw["op_off"], on["op_off"]) and np.array_equal(nw["ops"][:nw["n_ops"]], on["ops"])
    al.close()
