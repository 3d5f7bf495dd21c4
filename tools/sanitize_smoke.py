"""Small end-to-end pass over every kernel for `compute-sanitizer --tool memcheck` (run under gpurun)."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import __graft_entry__ as ge
import oracle_lib as ol

pkg = ge.load()
rng = np.random.default_rng(1)
for ml, k, n in ((128, 3, 3000), (192, 9, 1500), (256, 50, 300), (1024, 24, 200), (512, 50, 100)):
    W = ml // 64
    al = pkg.Aligner(k=k, max_length=ml)
    b = ol.random_batch(rng, n, 0, ml, 0.25, 0.3, 50)
    out = al.greedy_batch(b.reads, b.read_off, b.refs, b.ref_off)
    r = ol.oracle_greedy(b, ol.make_params(k, 1, 1, 1, W), threads=8)
    assert out["cost"].tolist() == r["cost"].tolist()
    p0, p1 = al.pack2bit_batch(b.reads, b.read_off)
    if ml <= 256:
        res = al.benchmark_batch(b.reads, b.read_off, b.refs, b.ref_off, want_penalties=True)
        nw = ol.oracle_nw(b, 1, 1, 1, threads=8)
        assert res["nw_penalty"].tolist() == nw["penalty"].tolist()
        assert res["counters"].tolist() == [int(v) for v in ol.oracle_metrics(b, r, nw)]
        al2 = pkg.Aligner(k=k, x=2, o=3, e=1, max_length=ml)
        g2 = al2.greedy_batch(b.reads, b.read_off, b.refs, b.ref_off, want_ops=False)["cost"]
        o2 = al2.nw_batch(b.reads, b.read_off, b.refs, b.ref_off, bound=g2)
        assert o2["penalty"].tolist() == ol.oracle_nw(b, 2, 3, 1, threads=8)["penalty"].tolist()
        al2.close()
    cfg = pkg.gen_config(5, min(ml - 28, 1000), 0.1)
    al.generate(cfg, 0, 500)
    al.close()
    print("ok", ml, k, flush=True)
# the widened band (gasma_set_correction): thread kernel and warp kernel
for ml, k, md, n in ((128, 3, 12, 2000), (192, 7, 20, 800), (512, 9, 16, 100)):
    W = ml // 64
    al = pkg.Aligner(k=k, max_length=ml)
    al.set_correction(True, md)
    b = ol.random_batch(rng, n, 0, ml, 0.2, 0.35, md)
    out = al.greedy_batch(b.reads, b.read_off, b.refs, b.ref_off)
    r = ol.oracle_greedy(b, ol.make_params(k, 1, 1, 1, W, correction=1), threads=8)
    assert out["cost"].tolist() == r["cost"].tolist()
    assert out["ops"][:out["n_ops"]].tolist() == r["ops"].tolist()
    al.close()
    print("ok widened band", ml, k, md, flush=True)
# the pipelined host call
al = pkg.Aligner()
b = ol.random_batch(rng, 66000, 0, 128, 0.2, 0.3, 50)
out = al.greedy_batch(b.reads, b.read_off, b.refs, b.ref_off)
assert out["cost"].tolist() == ol.oracle_greedy(b, ol.make_params(), threads=8)["cost"].tolist()
print("sanitize smoke ok")
