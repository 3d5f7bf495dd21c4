"""gasma_greedy_batch_dev_noalloc (include/gasma_b200.h): the greedy step whose scratch is a static __device__ array
of the library, so the call itself allocates nothing.  Collected last on purpose (file name).
Checked bit for bit against the existing entry point gasma_greedy_batch_dev, two calls in a row (the
second one finds the first one's contents in the shared array), and against the CPU oracle."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _dev_batch(torch, b):
    dev = torch.device("cuda", 0)
    pad = lambda a: np.concatenate([a, np.zeros(16, np.uint8)])
    return (torch.from_numpy(pad(b.reads)).to(dev), torch.from_numpy(b.read_off.view(np.int32)).to(dev),
            torch.from_numpy(pad(b.refs)).to(dev), torch.from_numpy(b.ref_off.view(np.int32)).to(dev))


def _run(torch, al, fn, d, n, cap):
    dev = torch.device("cuda", 0)
    cost = torch.full((n,), -7, dtype=torch.int32, device=dev)
    ops = torch.zeros(cap, dtype=torch.int16, device=dev)
    off = torch.full((n + 1,), -7, dtype=torch.int32, device=dev)
    torch.cuda.synchronize()
    free0 = torch.cuda.mem_get_info(0)[0]
    fn(d[0], d[1], d[2], d[3], n, cost, ops, off)
    al.synchronize()
    _run.taken = free0 - torch.cuda.mem_get_info(0)[0]   # device memory the call itself took
    off_h = off.cpu().numpy().view(np.uint32)
    return cost.cpu().numpy(), off_h, ops.cpu().numpy().view(np.uint16)[:int(off_h[n])]


@pytest.mark.parametrize("W,k", [(2, 3), (3, 5)])
def test_noalloc_entry_matches_existing_entry_twice(pkg, ol, W, k):
    import torch
    rng = np.random.default_rng(9100 + W)
    al = pkg.Aligner(k=k, max_length=64 * W)
    al.set_stream(torch.cuda.current_stream().cuda_stream)
    # two different batches back to back: the second call sees what the first left in the static array
    for n in (8192, 3001):
        b = ol.random_batch(rng, n, 1, 64 * W, 0.25, 0.3, 50)
        d = _dev_batch(torch, b)
        cap = n * (2 * 64 * W + 8)
        ref_cost, ref_off, ref_ops = _run(torch, al, al.greedy_batch_dev, d, n, cap)
        cost, off, ops = _run(torch, al, al.greedy_batch_dev_noalloc, d, n, cap)
        assert _run.taken < (1 << 20), "the no-alloc entry point took %d bytes of device memory" % _run.taken
        assert np.array_equal(cost, ref_cost)      # integer work: bit-exact
        assert np.array_equal(off, ref_off)
        assert np.array_equal(ops, ref_ops)
        r = ol.oracle_greedy(b, ol.make_params(k, 1, 1, 1, W, 1), threads=8)
        assert np.array_equal(cost, r["cost"])
        assert np.array_equal(off, r["op_off"])
        assert np.array_equal(ops, r["ops"])
    al.close()


def test_noalloc_entry_costs_only_and_limits(pkg, ol):
    import torch
    dev = torch.device("cuda", 0)
    rng = np.random.default_rng(9200)
    al = pkg.Aligner(k=3, max_length=128)
    al.set_stream(torch.cuda.current_stream().cuda_stream)
    b = ol.random_batch(rng, 2000, 1, 128, 0.2, 0.3, 50)
    d = _dev_batch(torch, b)
    cost = torch.empty(b.n, dtype=torch.int32, device=dev)
    al.greedy_batch_dev_noalloc(d[0], d[1], d[2], d[3], b.n, cost)
    al.synchronize()
    assert np.array_equal(cost.cpu().numpy(), ol.oracle_greedy(b, ol.make_params(), threads=8)["cost"])
    # sizes beyond the static array are rejected with a clear error, before anything is launched
    with pytest.raises(pkg.GasmaError) as ei:
        al.greedy_batch_dev_noalloc(d[0], d[1], d[2], d[3], 8193, cost)
    assert ei.value.code == -6 and "GASMA_NOALLOC_MAX_PAIRS" in str(ei.value)
    al.close()
    big = pkg.Aligner(k=3, max_length=512)
    with pytest.raises(pkg.GasmaError) as ei:
        big.greedy_batch_dev_noalloc(d[0], d[1], d[2], d[3], 10, cost)
    assert ei.value.code == -6 and "GASMA_NOALLOC_MAX_LENGTH" in str(ei.value)
    big.close()
