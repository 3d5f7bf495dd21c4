import os, sys, time
import numpy as np
sys.path.insert(0, '/root/repo')
import torch
import __graft_entry__ as ge
pkg = ge.load()
al = pkg.Aligner(k=3, max_length=192)
cfg = pkg.gen_config(0x47415342, 150, 0.10)
n = 1_000_000
reads, roff, refs, foff = al.generate(cfg, 0, n)
pin = lambda a: torch.from_numpy(np.ascontiguousarray(a)).pin_memory()
r0, r1 = al.pack2bit_stream(reads); f0, f1 = al.pack2bit_stream(refs)
tp = [pin(x.view(np.int16)) for x in (r0, r1, f0, f1)]
pk = [t.numpy().view(np.uint16) for t in tp]
to = [pin(roff.view(np.int32)), pin(foff.view(np.int32))]
ro, fo = to[0].numpy().view(np.uint32), to[1].numpy().view(np.uint32)
out = {"cost": torch.empty(n, dtype=torch.int32).pin_memory().numpy(), "ops": torch.empty(n * 24, dtype=torch.int16).pin_memory().numpy().view(np.uint16),
       "op_off": torch.empty(n + 1, dtype=torch.int32).pin_memory().numpy().view(np.uint32)}
for s in os.environ.get("SL", "8").split(","):
    os.environ["GASMA_SLICES"] = s
    a2 = pkg.Aligner(k=3, max_length=192)
    for i in range(4):
        t0 = time.perf_counter(); a2.greedy_batch_packed(pk[0], pk[1], ro, pk[2], pk[3], fo, out=out); dt = time.perf_counter() - t0
    print("slices", s, "call %.3f ms" % (dt * 1e3), flush=True)
    a2.close()
