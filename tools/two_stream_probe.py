#!/usr/bin/env python
"""Does the benchmark body (greedy + NW verifier + counters) gain from two batches in flight on two streams?

    python tools/two_stream_probe.py [pairs] [contexts...]

One context = the kernels of batch i+1 start when batch i is done.  Two contexts on two streams = the verifier of one
batch shares the GPU with the greedy kernel of the next: the tails of both persistent grids are filled.  Prints
alignments/s for 1, 2 and 3 contexts over the same resident batches (configs[2]: 100 bp, e = 0.20)."""
import importlib.util
import os
import sys
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
spec = importlib.util.spec_from_file_location("gasma_b200", os.path.join(ROOT, "approximate-string-matching_b200", "__init__.py"))
pkg = importlib.util.module_from_spec(spec)
sys.modules["gasma_b200"] = pkg
spec.loader.exec_module(pkg)


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
    ways = [int(v) for v in sys.argv[2:]] or [1, 2, 3]
    dev = torch.device("cuda:0")
    L, e, nb, rounds = 100, 0.20, 6, 4
    gen = pkg.Aligner(k=3, max_length=128)
    cfg = pkg.gen_config(0x47415341 + 2, L, e)
    batches = []
    for b in range(nb):
        d_reads = torch.empty(n * L + 16, dtype=torch.uint8, device=dev)
        d_refs = torch.empty(gen.generate_ref_capacity(cfg, n), dtype=torch.uint8, device=dev)
        d_roff = torch.empty(n + 1, dtype=torch.int32, device=dev)
        d_foff = torch.empty(n + 1, dtype=torch.int32, device=dev)
        gen.generate_dev(cfg, b * n, n, d_reads, d_roff, d_refs, d_foff)
        batches.append((d_reads, d_roff, d_refs, d_foff))
    gen.synchronize()
    ref = None
    for w in ways:
        als = [pkg.Aligner(k=3, max_length=128) for _ in range(w)]
        cnt = [torch.zeros(4, dtype=torch.int64, device=dev) for _ in range(w)]
        for i in range(2 * w):      # warm-up: allocations, first launches
            als[i % w].benchmark_batch_dev(*batches[i % nb], n, cnt[i % w])
        for a in als:
            a.synchronize()
        for c in cnt:
            c.zero_()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for i in range(nb * rounds):
            als[i % w].benchmark_batch_dev(*batches[i % nb], n, cnt[i % w])
        for a in als:
            a.synchronize()
        dt = time.perf_counter() - t0
        total = sum(c for c in cnt).tolist()
        if ref is None:
            ref = total
        print("%d context(s): %.4f ms per batch, %.1f M alignments/s, counters %s %s" %
              (w, 1e3 * dt / (nb * rounds), nb * rounds * n / dt / 1e6, total, "== 1 context" if total == ref else "DIFFER"))
        for a in als:
            a.close()


def greedy_only():
    """The same question for the greedy-only step (headline workload: 150 bp, e = 0.10, CIGAR CSR out)."""
    n = 1_000_000
    dev = torch.device("cuda:0")
    L, e, nb, rounds = 150, 0.10, 4, 6
    gen = pkg.Aligner(k=3, max_length=192)
    cfg = pkg.gen_config(0x47415341 + 1, L, e)
    batches = []
    for b in range(nb):
        d_reads = torch.empty(n * L + 16, dtype=torch.uint8, device=dev)
        d_refs = torch.empty(gen.generate_ref_capacity(cfg, n), dtype=torch.uint8, device=dev)
        d_roff = torch.empty(n + 1, dtype=torch.int32, device=dev)
        d_foff = torch.empty(n + 1, dtype=torch.int32, device=dev)
        gen.generate_dev(cfg, b * n, n, d_reads, d_roff, d_refs, d_foff)
        batches.append((d_reads, d_roff, d_refs, d_foff))
    gen.synchronize()
    for w in (1, 2, 3):
        als = [pkg.Aligner(k=3, max_length=192) for _ in range(w)]
        outs = [(torch.empty(n, dtype=torch.int32, device=dev), torch.empty(n * 104, dtype=torch.int16, device=dev),
                 torch.empty(n + 1, dtype=torch.int32, device=dev)) for _ in range(w)]
        for i in range(2 * w):
            als[i % w].greedy_batch_dev(*batches[i % nb], n, *outs[i % w])
        for a in als:
            a.synchronize()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for i in range(nb * rounds):
            als[i % w].greedy_batch_dev(*batches[i % nb], n, *outs[i % w])
        for a in als:
            a.synchronize()
        dt = time.perf_counter() - t0
        print("greedy only, %d context(s): %.4f ms per batch, %.1f M alignments/s" % (w, 1e3 * dt / (nb * rounds), nb * rounds * n / dt / 1e6))
        for a in als:
            a.close()


if __name__ == "__main__":
    if len(sys.argv) > 1 and sys.argv[1] == "greedy":
        greedy_only()
    else:
        main()
