import torch, time
n = 1 << 30
h = torch.empty(n, dtype=torch.uint8).pin_memory()
d = torch.empty(n, dtype=torch.uint8, device="cuda")
for sz in (1 << 30, 256 << 20, 64 << 20, 16 << 20):
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    reps = (1 << 31) // sz
    e0.record()
    for i in range(reps):
        d[:sz].copy_(h[:sz], non_blocking=True)
    e1.record(); torch.cuda.synchronize()
    print("H2D chunk %4d MiB: %.1f GB/s" % (sz >> 20, reps * sz / e0.elapsed_time(e1) / 1e6))
e0.record()
for i in range(2):
    h.copy_(d, non_blocking=True)
e1.record(); torch.cuda.synchronize()
print("D2H 1 GiB: %.1f GB/s" % (2 * n / e0.elapsed_time(e1) / 1e6))
s2 = torch.cuda.Stream()
h2 = torch.empty(n, dtype=torch.uint8).pin_memory(); d2 = torch.empty(n, dtype=torch.uint8, device="cuda")
torch.cuda.synchronize(); t0 = time.perf_counter()
for i in range(2):
    d.copy_(h, non_blocking=True)
    with torch.cuda.stream(s2):
        h2.copy_(d2, non_blocking=True)
torch.cuda.synchronize(); dt = time.perf_counter() - t0
print("bidirectional: %.1f GB/s each way" % (2 * n / dt / 1e9))
