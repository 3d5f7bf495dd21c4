"""Greedy kernel time per 1 M-pair batch as a function of the batch's first pair index (same process, same buffers)."""
import sys
import numpy as np
sys.path.insert(0, '/root/repo')
import torch
import __graft_entry__ as ge
pkg = ge.load()
L, e, n = 150, 0.10, 1_000_000
al = pkg.Aligner(k=3, max_length=192)
st = torch.cuda.current_stream(); al.set_stream(st.cuda_stream)
cfg = pkg.gen_config(0x47415341 + 1, L, e)
dev = torch.device("cuda", 0)
d_reads = torch.empty(n * L + 16, dtype=torch.uint8, device=dev)
d_refs = torch.empty(al.generate_ref_capacity(cfg, n), dtype=torch.uint8, device=dev)
d_roff = torch.empty(n + 1, dtype=torch.int32, device=dev); d_foff = torch.empty(n + 1, dtype=torch.int32, device=dev)
d_cost = torch.empty(n, dtype=torch.int32, device=dev); d_ops = torch.empty(n * 104, dtype=torch.int16, device=dev)
d_ooff = torch.empty(n + 1, dtype=torch.int32, device=dev)
for first in [0, 1, 2, 3, 4, 5, 6, 7, 0, 2, 15, 100]:
    al.generate_dev(cfg, first * n, n, d_reads, d_roff, d_refs, d_foff)
    for _ in range(3):
        al.greedy_batch_dev(d_reads, d_roff, d_refs, d_foff, n, d_cost, d_ops, d_ooff)
    al.synchronize(); al.kernel_time_ms(0); al.last_step_count()
    for _ in range(10):
        al.greedy_batch_dev(d_reads, d_roff, d_refs, d_foff, n, d_cost, d_ops, d_ooff)
    al.synchronize()
    ms, k = al.kernel_time_ms(0)
    steps = al.last_step_count() / 10
    print("first %3d M: kernel %.4f ms  steps/pair %.4f  ops %d  cost sum %d  ref bytes %d" % (
        first, ms / k, steps / n, int(d_ooff[-1].item()), int(d_cost.sum(dtype=torch.int64).item()), int(d_foff[-1].item())), flush=True)
