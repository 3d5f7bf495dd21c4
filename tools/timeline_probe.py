#!/usr/bin/env python
"""Where the time of ONE greedy launch goes: per-warp timeline (start, queue seen dry, last hand-out, exit).

    tools/build_variant.sh timeline -DGASMA_TIMELINE=1
    GASMA_B200_LIB=$PWD/_build/ab/timeline.so python tools/timeline_probe.py [pairs [length error max_length]]

The instrumented build (kernels.cuh, GASMA_TIMELINE: four globaltimer stamps per warp in a __device__ array, read back
through gasma_debug_timeline) is for this probe only; the shipped library is built without it."""
import ctypes as C, importlib.util, os, sys
import numpy as np, torch
ROOT = "/root/repo" if os.path.exists("/root/repo/bench.py") else os.getcwd()
spec = importlib.util.spec_from_file_location("gasma_b200", os.path.join(ROOT, "approximate-string-matching_b200", "__init__.py"))
pkg = importlib.util.module_from_spec(spec); sys.modules["gasma_b200"] = pkg; spec.loader.exec_module(pkg)
n = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
LEN = int(sys.argv[2]) if len(sys.argv) > 2 else 150
ERR = float(sys.argv[3]) if len(sys.argv) > 3 else 0.10
ML = int(sys.argv[4]) if len(sys.argv) > 4 else 192
dev = torch.device("cuda:0")
al = pkg.Aligner(k=3, max_length=ML)
cfg = pkg.gen_config(0x47415341 + 1, LEN, ERR)
bs = []
for b in range(2):
    d_reads = torch.empty(n * LEN + 16, dtype=torch.uint8, device=dev)
    d_refs = torch.empty(al.generate_ref_capacity(cfg, n), dtype=torch.uint8, device=dev)
    d_roff = torch.empty(n + 1, dtype=torch.int32, device=dev); d_foff = torch.empty(n + 1, dtype=torch.int32, device=dev)
    al.generate_dev(cfg, b * n, n, d_reads, d_roff, d_refs, d_foff); bs.append((d_reads, d_roff, d_refs, d_foff))
d_cost = torch.empty(n, dtype=torch.int32, device=dev); d_ops = torch.empty(n * 104, dtype=torch.int16, device=dev); d_ooff = torch.empty(n + 1, dtype=torch.int32, device=dev)
for i in range(4):
    al.greedy_batch_dev(*bs[i % 2], n, d_cost, d_ops, d_ooff)
al.synchronize()
L = pkg.lib()
out = np.zeros((4, 8192), np.uint64)
L.gasma_debug_timeline.argtypes = [C.c_void_p]
print("rc", L.gasma_debug_timeline(out.ctypes.data))
start, dry, last, ex = [out[i].astype(np.int64) for i in range(4)]
m = ex > 0
t0 = start[m].min()
print("warps", int(m.sum()), "kernel span us %.1f" % ((ex[m].max() - t0) / 1e3))
def pct(x, name):
    x = (x - t0) / 1e3
    print("%-22s min %.1f p10 %.1f p50 %.1f p90 %.1f p99 %.1f max %.1f" % (name, x.min(), *np.percentile(x, [10, 50, 90, 99]), x.max()))
pct(start[m], "start")
pct(dry[m & (dry > 0)], "queue seen dry")
pct(last[m & (last > 0)], "last hand-out")
pct(ex[m], "exit")
end = ex[m].max()
idle = (end - ex[m]).sum() / 1e3
print("idle warp-us after exit: %.0f = %.1f %% of span x warps" % (idle, 100 * idle / (((end - t0) / 1e3) * m.sum())))
