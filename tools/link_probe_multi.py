"""Host-to-device bandwidth per GPU with 1, 2, 4, 8 GPUs copying AT THE SAME TIME (one process, one thread and one
stream per GPU, pinned 256 MiB host buffers, CUDA-event timing per GPU).  Names the shared hop behind the e2e scaling:
per-link PCIe 5 x16 is ~55 GB/s; what every GPU gets when all copy together is what the host side can feed.

    python tools/link_probe_multi.py > gpurun_out/link_table.txt
"""
import threading
import time

import torch

G = torch.cuda.device_count()
SZ = 256 << 20
REPS = 16
hosts = [torch.empty(SZ, dtype=torch.uint8).pin_memory() for _ in range(G)]
devs = [torch.empty(SZ, dtype=torch.uint8, device="cuda:%d" % g) for g in range(G)]
for h in hosts:
    h.fill_(1)          # touch the pages


def worker(g, direction, barrier, out):
    torch.cuda.set_device(g)
    st = torch.cuda.Stream(device=g)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with torch.cuda.stream(st):
        for _ in range(2):
            devs[g].copy_(hosts[g], non_blocking=True)
        st.synchronize()
        barrier.wait()
        e0.record(st)
        for _ in range(REPS):
            if direction == "h2d":
                devs[g].copy_(hosts[g], non_blocking=True)
            else:
                hosts[g].copy_(devs[g], non_blocking=True)
        e1.record(st)
        st.synchronize()
    out[g] = REPS * SZ / e0.elapsed_time(e1) / 1e6


print("GPUs visible: %d, host threads: %d" % (G, torch.get_num_threads()))
for direction in ("h2d", "d2h"):
    for active in (1, 2, 4, 8):
        if active > G:
            continue
        out = {}
        barrier = threading.Barrier(active)
        ts = [threading.Thread(target=worker, args=(g, direction, barrier, out)) for g in range(active)]
        t0 = time.perf_counter()
        for t in ts:
            t.start()
        for t in ts:
            t.join()
        per = [out[g] for g in range(active)]
        print("%s, %d GPUs at once: per GPU %s GB/s | sum %.1f GB/s | min %.1f" % (
            direction, active, " ".join("%.1f" % v for v in per), sum(per), min(per)))
