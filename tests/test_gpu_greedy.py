"""GPU parity tests (run with -m gpu on a B200): the CUDA path through the C ABI against the golden
vectors of the unmodified reference and against the oracle on seeded inputs.  Bit-exact or fail."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _batch(ol, pairs):
    return ol.Batch.from_pairs([(a.encode("latin1"), b.encode("latin1")) for a, b in pairs])


def _run(pkg, al, b):
    out = al.greedy_batch(b.reads, b.read_off, b.refs, b.ref_off)
    text = [pkg.render_ops(out["ops"][out["op_off"][i]:out["op_off"][i + 1]]) for i in range(b.n)]
    return out, text


def test_gpu_matches_golden(pkg, ol, golden):
    aligners = {}
    for s in golden("greedy_ref.json")["sets"]:
        ps = s["params"]
        key = (ps["k"], ps["x"], ps["o"], ps["e"], tuple(ps["probs"]))
        if key not in aligners:
            aligners[key] = pkg.Aligner(k=ps["k"], x=ps["x"], o=ps["o"], e=ps["e"], match_prob=ps["probs"][0],
                                        mismatch_prob=ps["probs"][1], indel_prob=ps["probs"][2])
        b = _batch(ol, s["pairs"])
        out, text = _run(pkg, aligners[key], b)
        assert out["cost"].tolist() == s["cost"], s["name"]
        assert text == s["cigar"], s["name"]


@pytest.mark.parametrize("W", [1, 2, 3, 4])
def test_gpu_matches_oracle_all_widths(pkg, ol, W):
    rng = np.random.default_rng(300 + W)
    for (k, x, o, e, g) in [(3, 1, 1, 1, 0), (3, 1, 1, 1, 1), (9, 2, 3, 1, 1), (25, 1, 2, 1, 1)]:
        b = ol.random_batch(rng, 5000, 1, 64 * W, 0.3, 0.3, 50)
        al = pkg.Aligner(k=k, x=x, o=o, e=e, max_length=64 * W, alignment_type=1 - g)
        out, _ = _run(pkg, al, b)
        r = ol.oracle_greedy(b, ol.make_params(k, x, o, e, W, g), threads=8)
        assert out["cost"].tolist() == r["cost"].tolist()
        assert out["op_off"].tolist() == r["op_off"].tolist()
        assert out["ops"][:out["n_ops"]].tolist() == r["ops"].tolist()
        al.close()


def test_gpu_pack2bit(pkg, ol, golden):
    vec = golden("convert2bit_ref.json")["vectors"]
    seqs = np.frombuffer("".join(v["in"] for v in vec).encode("latin1"), np.uint8)
    off = (np.arange(len(vec) + 1) * 128).astype(np.uint32)
    al = pkg.Aligner()
    p0, p1 = al.pack2bit_batch(seqs, off)
    for i, v in enumerate(vec):
        assert p0[i].view(np.uint8).tolist() == v["bits0"]
        assert p1[i].view(np.uint8).tolist() == v["bits1"]
    rng = np.random.default_rng(6)
    for W in (1, 2, 3, 4):
        a = pkg.Aligner(max_length=64 * W)
        lens = rng.integers(0, 64 * W + 1, 3000)
        off = np.zeros(len(lens) + 1, np.uint32)
        off[1:] = np.cumsum(lens)
        seqs = rng.choice(np.frombuffer(b"ACGTNacgt\0\xff", np.uint8), int(off[-1])).astype(np.uint8)
        p0, p1 = a.pack2bit_batch(seqs, off)
        for i in range(0, len(lens), 7):
            e0, e1 = ol.oracle_pack(seqs[off[i]:off[i + 1]].tobytes(), W)
            assert p0[i].tolist() == e0.tolist() and p1[i].tolist() == e1.tolist()


def test_gpu_edge_cases(pkg, ol):
    al = pkg.Aligner()
    # empty batch, empty strings, single pair, costs only
    out = al.greedy_batch(np.zeros(0, np.uint8), np.zeros(1, np.uint32), np.zeros(0, np.uint8), np.zeros(1, np.uint32))
    assert out["n_ops"] == 0
    b = _batch(ol, [("", ""), ("A", ""), ("", "A"), ("ACGT", "ACGT")])
    out, text = _run(pkg, al, b)
    assert out["cost"].tolist() == [0, 1, 1, 0] and text == ["", "1I", "1D", "4M"]
    out = al.greedy_batch(b.reads, b.read_off, b.refs, b.ref_off, want_ops=False)
    assert out["cost"].tolist() == [0, 1, 1, 0]
    # errors: too long, |n-m| > MAX_K, ops capacity
    with pytest.raises(pkg.GasmaError) as ei:
        bb = _batch(ol, [("A" * 129, "A" * 128)])
        al.greedy_batch(bb.reads, bb.read_off, bb.refs, bb.ref_off)
    assert ei.value.code == -2
    with pytest.raises(pkg.GasmaError) as ei:
        bb = _batch(ol, [("A" * 10, "A" * 61)])
        al.greedy_batch(bb.reads, bb.read_off, bb.refs, bb.ref_off)
    assert ei.value.code == -3
    with pytest.raises(pkg.GasmaError) as ei:
        al.greedy_batch(b.reads, b.read_off, b.refs, b.ref_off, ops_cap=1)
    assert ei.value.code == -4
    # the context is still usable after an error, and the band can be changed between batches
    al.set_band(5)
    out, text = _run(pkg, al, b)
    assert text == ["", "1I", "1D", "4M"]


def test_gpu_generator_and_device_path(pkg, ol):
    """Device generator == CPU twin; resident (_dev) path == host path; 1M-pair property checks."""
    import torch
    al = pkg.Aligner()
    for mode, e, prof in [(0, 0.10, None), (1, 0.20, None), (2, 0.0, (0.02452, 0.000468, 0.000553))]:
        cfg = pkg.gen_config(0x47415341, 100, e, mode=mode, profile=prof)
        ocfg = ol.gen_config(0x47415341, 100, e, mode=mode, profile=prof)
        n = 3000
        d_reads = torch.empty(n * 100 + 16, dtype=torch.uint8, device="cuda")
        d_refs = torch.empty(al.generate_ref_capacity(cfg, n), dtype=torch.uint8, device="cuda")
        d_roff = torch.empty(n + 1, dtype=torch.int32, device="cuda")
        d_foff = torch.empty(n + 1, dtype=torch.int32, device="cuda")
        al.generate_dev(cfg, 12345, n, d_reads, d_roff, d_refs, d_foff)
        al.synchronize()
        ref = ol.generate(ocfg, 12345, n)
        foff = d_foff.cpu().numpy().view(np.uint32)
        assert foff.tolist() == ref.ref_off.tolist()
        assert d_roff.cpu().numpy().view(np.uint32).tolist() == ref.read_off.tolist()
        assert d_reads.cpu().numpy()[:n * 100].tobytes() == ref.reads.tobytes()
        assert d_refs.cpu().numpy()[:foff[-1]].tobytes() == ref.refs.tobytes()
        # resident path
        d_cost = torch.empty(n, dtype=torch.int32, device="cuda")
        d_ops = torch.empty(n * 64, dtype=torch.int16, device="cuda")
        d_ooff = torch.empty(n + 1, dtype=torch.int32, device="cuda")
        al.greedy_batch_dev(d_reads, d_roff, d_refs, d_foff, n, d_cost, d_ops, d_ooff)
        al.synchronize()
        r = ol.oracle_greedy(ref, ol.make_params())
        assert d_cost.cpu().numpy().tolist() == r["cost"].tolist()
        ooff = d_ooff.cpu().numpy().view(np.uint32)
        assert ooff.tolist() == r["op_off"].tolist()
        assert d_ops.cpu().numpy().view(np.uint16)[:ooff[-1]].tolist() == r["ops"].tolist()


def test_gpu_full_size_properties(pkg, ol):
    """BASELINE config sizes (1M pairs): size-independent properties + sampled oracle parity."""
    import torch
    al = pkg.Aligner()
    n = 1_000_000
    cfg = pkg.gen_config(0x47415342, 100, 0.10)
    d_reads = torch.empty(n * 100 + 16, dtype=torch.uint8, device="cuda")
    d_refs = torch.empty(al.generate_ref_capacity(cfg, n), dtype=torch.uint8, device="cuda")
    d_roff = torch.empty(n + 1, dtype=torch.int32, device="cuda")
    d_foff = torch.empty(n + 1, dtype=torch.int32, device="cuda")
    al.generate_dev(cfg, 0, n, d_reads, d_roff, d_refs, d_foff)
    d_cost = torch.empty(n, dtype=torch.int32, device="cuda")
    d_ops = torch.empty(n * 32, dtype=torch.int16, device="cuda")
    d_ooff = torch.empty(n + 1, dtype=torch.int32, device="cuda")
    al.greedy_batch_dev(d_reads, d_roff, d_refs, d_foff, n, d_cost, d_ops, d_ooff)
    al.synchronize()
    cost = d_cost.cpu().numpy()
    ooff = d_ooff.cpu().numpy().view(np.uint32).astype(np.int64)
    ops = d_ops.cpu().numpy().view(np.uint16)[:ooff[-1]]
    assert cost.min() >= 0 and cost.max() <= 60
    assert (np.diff(ooff) >= 1).all()                       # sorted CSR, every 100 bp pair has >= 1 op
    # idempotence: a second run gives identical bytes
    d_cost2 = torch.empty_like(d_cost); d_ops2 = torch.empty_like(d_ops); d_ooff2 = torch.empty_like(d_ooff)
    al.greedy_batch_dev(d_reads, d_roff, d_refs, d_foff, n, d_cost2, d_ops2, d_ooff2)
    al.synchronize()
    assert torch.equal(d_cost, d_cost2) and torch.equal(d_ooff, d_ooff2)
    assert torch.equal(d_ops[:ooff[-1]], d_ops2[:ooff[-1]])
    # CIGAR consumes the strings consistently: sum(M)+sum(I) and sum(M)+sum(D) per pair are bounded by lengths,
    # and the penalty is at least the number of indel ops (x=o=e=1)
    code = ops & 7
    ln = (ops >> 3).astype(np.int64)
    pair_of = np.repeat(np.arange(n), np.diff(ooff))
    indel = np.bincount(pair_of, weights=(code != 0) * ln, minlength=n)
    assert (cost >= indel).all()
    # sampled parity against the oracle: first / middle / last 2000 pairs
    ocfg = ol.gen_config(0x47415342, 100, 0.10)
    for lo in (0, n // 2, n - 2000):
        ref = ol.generate(ocfg, lo, 2000)
        r = ol.oracle_greedy(ref, ol.make_params(), threads=8)
        assert cost[lo:lo + 2000].tolist() == r["cost"].tolist()
        seg = ops[ooff[lo]:ooff[lo + 2000]]
        assert seg.tolist() == r["ops"].tolist()


def test_gpu_host_call_pipelined_slices(pkg, ol):
    """Host-pointer call on a batch large enough to be pipelined in slices (H2D / kernels / D2H overlapped): one CSR
    array across the slices, bit-equal to the oracle and to the resident (_dev) path; capacity errors still reported."""
    import torch
    al = pkg.Aligner()
    rng = np.random.default_rng(4242)
    b = ol.random_batch(rng, 70_001, 0, 128, 0.25, 0.3, 50)     # ragged lengths, empty strings, odd pair count
    out = al.greedy_batch(b.reads, b.read_off, b.refs, b.ref_off)
    r = ol.oracle_greedy(b, ol.make_params(), threads=8)
    assert out["cost"].tolist() == r["cost"].tolist()
    assert out["op_off"].tolist() == r["op_off"].tolist()
    assert out["ops"][:out["n_ops"]].tolist() == r["ops"].tolist()
    # costs only
    c2 = al.greedy_batch(b.reads, b.read_off, b.refs, b.ref_off, want_ops=False)["cost"]
    assert c2.tolist() == r["cost"].tolist()
    # too small an ops buffer: GASMA_E_CAPACITY with op_off[n] = size needed
    small = {"cost": np.empty(b.n, np.int32), "ops": np.empty(1000, np.uint16), "op_off": np.empty(b.n + 1, np.uint32)}
    with pytest.raises(pkg.GasmaError) as ei:
        al.greedy_batch(b.reads, b.read_off, b.refs, b.ref_off, ops_cap=1000, out=small)
    assert ei.value.code == -4 and int(small["op_off"][b.n]) == len(r["ops"])
    assert small["cost"].tolist() == r["cost"].tolist()
    # 1M-pair synthetic batch through the host call == the resident path
    cfg = pkg.gen_config(0x47415360, 100, 0.10)
    reads, roff, refs, foff = al.generate(cfg, 0, 1_000_000)
    h = al.greedy_batch(reads, roff, refs, foff)
    d = [torch.from_numpy(x.view(np.uint8) if x.dtype == np.uint8 else x.view(np.int32)).cuda() for x in (reads, roff, refs, foff)]
    d_reads = torch.zeros(len(reads) + 16, dtype=torch.uint8, device="cuda"); d_reads[:len(reads)] = d[0]
    d_refs = torch.zeros(len(refs) + 16, dtype=torch.uint8, device="cuda"); d_refs[:len(refs)] = d[2]
    d_cost = torch.empty(1_000_000, dtype=torch.int32, device="cuda")
    d_ops = torch.empty(h["n_ops"] + 8, dtype=torch.int16, device="cuda")
    d_ooff = torch.empty(1_000_001, dtype=torch.int32, device="cuda")
    al.greedy_batch_dev(d_reads, d[1], d_refs, d[3], 1_000_000, d_cost, d_ops, d_ooff)
    al.synchronize()
    assert np.array_equal(h["cost"], d_cost.cpu().numpy())
    assert np.array_equal(h["op_off"], d_ooff.cpu().numpy().view(np.uint32))
    assert np.array_equal(h["ops"][:h["n_ops"]], d_ops.cpu().numpy().view(np.uint16)[:h["n_ops"]])
    al.close()


@pytest.mark.parametrize("W,k", [(8, 3), (8, 24), (16, 3), (16, 24), (16, 50)])
def test_gpu_long_reads_warp_per_alignment(pkg, ol, W, k):
    """max_length 512 / 1024 (BASELINE configs[4]: 1 kbp reads, wide band, multi-word bitmasks): the
    one-alignment-per-warp kernel == the W-word oracle on cost and CIGAR ops, for narrow and wide bands."""
    rng = np.random.default_rng(8000 + 10 * W + k)
    al = pkg.Aligner(k=k, max_length=64 * W)
    for (x, o, e, n, max_err) in [(1, 1, 1, 1500, 0.2), (2, 3, 1, 600, 0.3)]:
        al2 = al if (x, o, e) == (1, 1, 1) else pkg.Aligner(k=k, x=x, o=o, e=e, max_length=64 * W)
        b = ol.random_batch(rng, n, 0, 64 * W, max_err, 0.3, 50)
        r = ol.oracle_greedy(b, ol.make_params(k, x, o, e, W), threads=8)
        out = al2.greedy_batch(b.reads, b.read_off, b.refs, b.ref_off)
        assert out["cost"].tolist() == r["cost"].tolist()
        assert out["op_off"].tolist() == r["op_off"].tolist()
        assert out["ops"][:out["n_ops"]].tolist() == r["ops"].tolist()
        if al2 is not al:
            al2.close()
    # the named long-read workload: 1 kbp reads at 15 % error (exact edit count), generated on the device
    if W == 16:
        cfg = pkg.gen_config(0x47415370, 1000, 0.15)
        reads, roff, refs, foff = al.generate(cfg, 0, 3000)
        b = ol.Batch(reads, roff, refs, foff)
        r = ol.oracle_greedy(b, ol.make_params(k, 1, 1, 1, W), threads=8)
        out = al.greedy_batch(b.reads, b.read_off, b.refs, b.ref_off)
        assert out["cost"].tolist() == r["cost"].tolist()
        assert out["ops"][:out["n_ops"]].tolist() == r["ops"].tolist()
    # packing at this width
    seqs = rng.choice(np.frombuffer(b"ACGTNacgt", np.uint8), 64 * W * 40).astype(np.uint8)
    off = (np.arange(41) * 64 * W).astype(np.uint32)
    p0, p1 = al.pack2bit_batch(seqs, off)
    for i in range(0, 40, 7):
        e0, e1 = ol.oracle_pack(seqs[off[i]:off[i + 1]].tobytes(), W)
        assert p0[i].tolist() == e0.tolist() and p1[i].tolist() == e1.tolist()
    with pytest.raises(pkg.GasmaError):   # the NW verifier stops at 256 characters
        al.nw_batch(b.reads, b.read_off, b.refs, b.ref_off)
    al.close()


def test_gpu_widened_band_matches_golden(pkg, ol, golden):
    """gasma_set_correction against the unmodified reference compiled with -DCORRECTION (golden fixture)."""
    aligners = {}
    for s in golden("greedy_ref_correction.json")["sets"]:
        ps = s["params"]
        key = (ps["k"], ps["x"], ps["o"], ps["e"], tuple(ps["probs"]))
        if key not in aligners:
            aligners[key] = pkg.Aligner(k=ps["k"], x=ps["x"], o=ps["o"], e=ps["e"], match_prob=ps["probs"][0],
                                        mismatch_prob=ps["probs"][1], indel_prob=ps["probs"][2])
            aligners[key].set_correction(True, ps["max_delta"])
        b = _batch(ol, s["pairs"])
        out, text = _run(pkg, aligners[key], b)
        assert out["cost"].tolist() == s["cost"], s["name"]
        assert text == s["cigar"], s["name"]


@pytest.mark.parametrize("W", [2, 3, 8, 16])
def test_gpu_widened_band_matches_oracle(pkg, ol, W):
    """Thread kernel (W <= 4) and warp kernel (W = 8, 16) with the widened band, against the oracle in that mode;
    switching it off again restores lanes -k..k; pairs beyond the lane slots fail the batch."""
    rng = np.random.default_rng(700 + W)
    n = 4000 if W <= 4 else 600
    for (k, x, o, e, g, md) in [(3, 1, 1, 1, 1, 12), (7, 2, 3, 1, 1, 20), (3, 1, 1, 1, 0, 9)]:
        b = ol.random_batch(rng, n, 1, 64 * W, 0.2 if W <= 4 else 0.1, 0.35, md)
        al = pkg.Aligner(k=k, x=x, o=o, e=e, max_length=64 * W, alignment_type=1 - g)
        al.set_correction(True, md)
        out, _ = _run(pkg, al, b)
        r = ol.oracle_greedy(b, ol.make_params(k, x, o, e, W, g, correction=1), threads=8)
        assert out["cost"].tolist() == r["cost"].tolist()
        assert out["op_off"].tolist() == r["op_off"].tolist()
        assert out["ops"][:out["n_ops"]].tolist() == r["ops"].tolist()
        al.set_correction(False)
        out, _ = _run(pkg, al, b)
        r = ol.oracle_greedy(b, ol.make_params(k, x, o, e, W, g), threads=8)
        assert out["cost"].tolist() == r["cost"].tolist()
        assert out["ops"][:out["n_ops"]].tolist() == r["ops"].tolist()
        al.close()
    al = pkg.Aligner(max_length=64 * W)
    al.set_correction(True, 4)
    b = _batch(ol, [("ACGT" * 5, "ACGT" * 7), ("ACGT", "ACGT")])
    with pytest.raises(pkg.GasmaError) as ei:
        al.greedy_batch(b.reads, b.read_off, b.refs, b.ref_off)
    assert ei.value.code == -3   # GASMA_E_BAND
    with pytest.raises(pkg.GasmaError):
        al.set_correction(True, 48)    # max_delta + k > MAX_K


@pytest.mark.parametrize("variant", [1, 2])
def test_gpu_mask_variants(pkg, ol, golden, variant):
    """gasma_set_mask_variant: the highway search on flip_short_hurdles(2) / flip_short_hurdles(1).flip_short_matches(1)
    (utils.h:200-240), against the unmodified reference's outputs for that call site and against the oracle."""
    for s in golden("greedy_ref_mask.json")["sets"]:
        if s["variant"] != variant:
            continue
        ps = s["params"]
        al = pkg.Aligner(k=ps["k"], x=ps["x"], o=ps["o"], e=ps["e"], match_prob=ps["probs"][0], mismatch_prob=ps["probs"][1],
                         indel_prob=ps["probs"][2])
        al.set_mask_variant(variant)
        out, text = _run(pkg, al, _batch(ol, s["pairs"]))
        assert out["cost"].tolist() == s["cost"], s["name"]
        assert text == s["cigar"], s["name"]
        al.close()
    rng = np.random.default_rng(660 + variant)
    for W in (1, 2, 3, 4):
        b = ol.random_batch(rng, 20000, 1, 64 * W, 0.3, 0.3, 50)
        al = pkg.Aligner(k=5, x=1, o=2, e=1, max_length=64 * W)
        al.set_mask_variant(variant)
        out, _ = _run(pkg, al, b)
        r = ol.oracle_greedy(b, ol.make_params(5, 1, 2, 1, W, 1, mask_variant=variant), threads=8)
        assert out["cost"].tolist() == r["cost"].tolist()
        assert out["op_off"].tolist() == r["op_off"].tolist()
        assert out["ops"][:out["n_ops"]].tolist() == r["ops"].tolist()
        al.set_mask_variant(0)      # and back: the default aligner again
        out0, _ = _run(pkg, al, b)
        assert out0["cost"].tolist() == ol.oracle_greedy(b, ol.make_params(5, 1, 2, 1, W, 1), threads=8)["cost"].tolist()
        al.close()
    with pytest.raises(pkg.GasmaError):
        pkg.Aligner(max_length=512).set_mask_variant(1)


@pytest.mark.parametrize("W", [1, 2, 3, 4])
def test_gpu_packed_input_equals_ascii_input(pkg, ol, W):
    """gasma_pack2bit_stream + gasma_greedy_batch_packed (2-bit streams in, 4x fewer bytes over the link) give exactly
    what gasma_greedy_batch gives on the ASCII form of the same batch -- small (one copy) and large (pipelined slices)
    batches, ragged lengths, non-ACGT bytes; and the streams themselves equal the oracle's packing bit for bit."""
    rng = np.random.default_rng(7300 + W)
    al = pkg.Aligner(k=4, max_length=64 * W)
    for n, alphabet in ((3000, b"ACGT"), (70000, b"ACGT"), (2000, b"ACGTNacgt")):
        b = ol.random_batch(rng, n, 0, 64 * W, 0.3, 0.3, 50, alphabet=alphabet)
        r0, r1 = al.pack2bit_stream(b.reads)
        f0, f1 = al.pack2bit_stream(b.refs)
        # stream bit i = plane bit of character i (checked against the oracle's per-string packing on a few strings)
        for i in rng.integers(0, b.n, 40):
            lo, hi = int(b.read_off[i]), int(b.read_off[i + 1])
            p0, p1 = ol.oracle_pack(b.reads[lo:hi].tobytes(), W)
            bits0 = np.unpackbits(r0.view(np.uint8), bitorder="little")[lo:hi]
            want0 = np.unpackbits(p0.view(np.uint8), bitorder="little")[:hi - lo]
            assert np.array_equal(bits0, want0)
            bits1 = np.unpackbits(r1.view(np.uint8), bitorder="little")[lo:hi]
            assert np.array_equal(bits1, np.unpackbits(p1.view(np.uint8), bitorder="little")[:hi - lo])
        a = al.greedy_batch(b.reads, b.read_off, b.refs, b.ref_off)
        p = al.greedy_batch_packed(r0, r1, b.read_off, f0, f1, b.ref_off)
        assert np.array_equal(a["cost"], p["cost"])
        assert np.array_equal(a["op_off"], p["op_off"])
        assert np.array_equal(a["ops"][:a["n_ops"]], p["ops"][:p["n_ops"]])
        r = ol.oracle_greedy(b, ol.make_params(4, 1, 1, 1, W, 1), threads=8)
        assert np.array_equal(p["cost"], r["cost"]) and np.array_equal(p["ops"][:p["n_ops"]], r["ops"])
    al.close()


def test_gpu_async_submit_wait_pair(pkg, ol):
    """gasma_greedy_submit / gasma_greedy_submit_packed / gasma_greedy_wait: two calls in flight give the synchronous
    calls' results bit for bit (small one-copy and large pipelined batches, ASCII and packed, different settings per
    round), a third submit is refused, unknown tickets are refused, errors of a submitted call arrive at its wait."""
    rng = np.random.default_rng(9100)
    al = pkg.Aligner(k=3, max_length=192)
    batches = [ol.random_batch(rng, n, 0, 192, 0.25, 0.3, 50) for n in (70000, 3000, 90000, 500)]
    sync = [al.greedy_batch(b.reads, b.read_off, b.refs, b.ref_off) for b in batches]
    ref0 = ol.oracle_greedy(batches[0], ol.make_params(3, 1, 1, 1, 3, 1), threads=8)
    assert np.array_equal(sync[0]["cost"], ref0["cost"]) and np.array_equal(sync[0]["ops"][:sync[0]["n_ops"]], ref0["ops"])

    def same(a, b):
        return (np.array_equal(a["cost"], b["cost"]) and np.array_equal(a["op_off"], b["op_off"])
                and np.array_equal(a["ops"][:a["n_ops"]], b["ops"][:b["n_ops"]]))

    # a rolling window of two calls over the four batches, three times round
    order = [0, 1, 2, 3] * 3
    pending = []
    for i in order:
        b = batches[i]
        if len(pending) == 2:
            t, j = pending.pop(0)
            assert same(al.greedy_wait(t), sync[j])
        pending.append((al.greedy_submit(b.reads, b.read_off, b.refs, b.ref_off), i))
    # both slots busy now: a third submit is refused, and nothing about the two in flight changes
    assert len(pending) == 2
    with pytest.raises(pkg.GasmaError) as ei:
        al.greedy_submit(batches[3].reads, batches[3].read_off, batches[3].refs, batches[3].ref_off)
    assert ei.value.code == -6 and "in flight" in str(ei.value)
    # a synchronous call on the parent may be mixed in while two are in flight
    assert same(al.greedy_batch(batches[1].reads, batches[1].read_off, batches[1].refs, batches[1].ref_off), sync[1])
    for t, j in pending:
        assert same(al.greedy_wait(t), sync[j])
    with pytest.raises(pkg.GasmaError) as ei:
        al.greedy_wait(12345)
    assert ei.value.code == -1

    # packed input, and the parent's settings at submit time apply (band 5, mask variant 1)
    al.set_band(5)
    al.set_mask_variant(1)
    b = batches[2]
    want = al.greedy_batch(b.reads, b.read_off, b.refs, b.ref_off)
    assert not np.array_equal(want["cost"], sync[2]["cost"])
    r0, r1 = al.pack2bit_stream(b.reads)
    f0, f1 = al.pack2bit_stream(b.refs)
    t1 = al.greedy_submit_packed(r0, r1, b.read_off, f0, f1, b.ref_off)
    t2 = al.greedy_submit(b.reads, b.read_off, b.refs, b.ref_off)
    assert same(al.greedy_wait(t2), want) and same(al.greedy_wait(t1), want)
    al.set_band(3)
    al.set_mask_variant(0)

    # an error inside a submitted call (a string longer than max_length) comes back from its wait, and the slot is free again
    bad = ol.Batch.from_pairs([(b"ACGT" * 60, b"ACGT" * 60)])       # 240 > 192
    t = al.greedy_submit(bad.reads, bad.read_off, bad.refs, bad.ref_off)
    with pytest.raises(pkg.GasmaError) as ei:
        al.greedy_wait(t)
    assert ei.value.code == -2
    t = al.greedy_submit(batches[1].reads, batches[1].read_off, batches[1].refs, batches[1].ref_off)
    assert same(al.greedy_wait(t), sync[1])
    # destroying a context with a call in flight waits for it
    t = al.greedy_submit(batches[0].reads, batches[0].read_off, batches[0].refs, batches[0].ref_off)
    al.close()


@pytest.mark.parametrize("W", [1, 3])
def test_gpu_end_game_boundaries(pkg, ol, W):
    """The greedy kernel's end game (greedy_coop.cuh) finishes a warp's last open pairs with one thread per lane when the
    pair has at most 32 lanes, and leaves them to their owner otherwise: band widths on both sides of that limit
    (k = 0, 15, 16), batches of 1 ... a few pools (where nearly every pair ends in the end game), GLOBAL and SEMI_GLOBAL,
    many-step pairs (high error rates), against the checker op for op."""
    rng = np.random.default_rng(4400 + W)
    for k in (0, 1, 15, 16):
        for g in (1, 0):
            al = pkg.Aligner(k=k, x=1, o=2, e=1, max_length=64 * W, alignment_type=1 - g)
            for n in (1, 2, 33, 130, 1500):
                b = ol.random_batch(rng, n, 1, 64 * W, 0.45, 0.4, 50)
                out, _ = _run(pkg, al, b)
                r = ol.oracle_greedy(b, ol.make_params(k, 1, 2, 1, W, g), threads=4)
                assert out["cost"].tolist() == r["cost"].tolist(), (k, g, n)
                assert out["op_off"].tolist() == r["op_off"].tolist(), (k, g, n)
                assert out["ops"][:out["n_ops"]].tolist() == r["ops"].tolist(), (k, g, n)
            al.close()
