#!/usr/bin/env python
"""bench.py -- alignments/s of the greedy hot path on B200 (and of the reference's CPU path beside it).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

A "step" is one pass of the hot path (pack -> lanes -> greedy traversal -> penalty + CIGAR CSR) over one batch
of synthetic pairs.  Default workload = BASELINE.json configs[1]: simulated 150 bp reads, error rate 0.10,
1M alignments per batch per GPU, greedy path (k=3, x=o=e=1, max_length 192 = 3 words).  `value` is measured
with the batch resident in HBM; `e2e` goes through the host-pointer C-ABI call with pinned host buffers
(H2D + kernels + D2H inside the timed region).  Ranks shard by pair index (weak scaling: every GPU gets its
own 1M-pair batches); the only collective is the final sum of the result counters.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

SEED = 0x47415341  # "GASA"
PROFILE_ILLUMINA = (0.02452, 0.000468, 0.000553)  # README.md:71-76


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--pairs", type=int, default=1_000_000, help="pairs per batch per GPU")
    ap.add_argument("--length", type=int, default=150)
    ap.add_argument("--error", type=float, default=0.10)
    ap.add_argument("--k", type=int, default=3)
    ap.add_argument("--profile", default="simulated", choices=["simulated", "illumina"],
                    help="simulated = exact edit budget (Dataset); illumina = per-base SRR611076 rates (BASELINE configs[3])")
    ap.add_argument("--batches", type=int, default=4, help="distinct resident batches per GPU cycled through (kernel time "
                    "varies +-3.5 %% between batches of identical total work: a tail effect of the dynamic chunk queue)")
    ap.add_argument("--cpu-sample", type=int, default=0, help="pairs in the CPU baseline sample (0 = auto)")
    ap.add_argument("--in-flight", type=int, default=3, help="batches in flight (contexts / streams) in the in_flight passes of the headline and full-path legs; 1 = off")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-full-path", action="store_true", help="skip the greedy + NW verifier + counters leg")
    ap.add_argument("--no-configs", action="store_true", help="skip the short legs for the other BASELINE.json configs")
    ap.add_argument("--no-parity", action="store_true", help="skip the oracle comparison of the timed outputs")
    return ap.parse_args()


def max_length_for(length, error):
    """Smallest supported max_length that holds the refs of this workload.  Only 4 % of the edits are indels
    (mismatch_rate 0.96), so a ref outgrows its read by a handful of bases, not by the whole edit budget; a pair that
    did exceed max_length would fail the run loudly (GASMA_E_LENGTH), never silently."""
    ne = int(np.ceil(np.float32(length) * np.float32(error)))
    need = length + (min(ne, max(8, int(np.ceil(ne * 0.04 * 4)))) if ne else 8)   # ne == 0: per-base profile, a few inserts
    for ml in (64, 128, 192, 256, 512, 1024):
        if need <= ml:
            return ml
    raise SystemExit("workload does not fit max_length 1024")


def gen_config_for(pkg, a, seed):
    if a.profile == "illumina":
        return pkg.gen_config(seed + 16, a.length, 0.0, mode=2, profile=PROFILE_ILLUMINA)
    return pkg.gen_config(seed, a.length, a.error)


def config_dict(a, ml):
    """The SAME dict from both arms (the driver compares them)."""
    return {"workload": workload_name(a), "max_length": ml, "length": a.length, "error_rate": a.error, "profile": a.profile,
            "pairs_per_batch_per_gpu": a.pairs, "k": a.k, "penalties": "x=o=e=1",
            "path": "greedy only: pack2bit + lanes + traversal + penalty + CIGAR (reset + run + get_cost + get_CIGAR)"}


def lib_sha(pkg):
    import hashlib
    h = hashlib.sha1()
    with open(pkg.LIB_PATH, "rb") as f:
        for blk in iter(lambda: f.read(1 << 20), b""):
            h.update(blk)
    return h.hexdigest()[:16]


def source_sha(pkg):
    """Hash of the sources in this tree (approximate-string-matching_b200/build.py:source_sha)."""
    import importlib.util
    path = os.path.join(os.path.dirname(os.path.abspath(pkg.__file__)), "build.py")
    spec = importlib.util.spec_from_file_location("gasma_b200_build", path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod.source_sha()


def ncu_figures(pkg, leg):
    """Per-launch DRAM bytes / executed instructions of the leg's dominant kernel from profiles/traffic.json -- only if
    they were taken from a build of the very sources the loaded library was built from (source-hash match); otherwise None.  Written by
    tools/update_traffic.py from a metrics-only ncu pass of this bench command (tools/profile_round.sh)."""
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            t = json.load(f)
        # the same code = the same source hash (nvcc output is not byte-reproducible: the .so's own SHA-1 changes with
        # every rebuild, so it only matches on the very file the capture ran on)
        in_tree = os.path.dirname(os.path.abspath(pkg.LIB_PATH)) == os.path.dirname(os.path.abspath(pkg.__file__))
        # (a library from elsewhere -- GASMA_B200_LIB, the A/B variants -- was not built from this tree's sources)
        if not (in_tree and t.get("source_sha1") == source_sha(pkg)) and t.get("library_sha1") != lib_sha(pkg):
            return None
        fig = dict(t["legs"].get(leg) or {})
        if not fig:
            return None
        fig["source_sha"] = t.get("source_sha1")
        fig["library_sha1_at_capture"] = t.get("library_sha1")
        return fig
    except Exception:
        return None


def attach_ncu(roof, fig, k_ms, sm_count, sm_mhz):
    """roofline.traffic + the issue-slot view of the same launch (instructions from ncu, time from the live run)."""
    if not fig:
        roof["traffic_note"] = "no ncu figures for these sources in profiles/traffic.json (source hash mismatch): traffic left null"
        return
    roof["traffic"] = fig["traffic_bytes"]
    roof["ncu"] = {k: fig[k] for k in ("kernel", "issue_active_pct", "alu_pipe_pct", "avg_active_threads", "warp_instructions",
                                       "dram_read_bytes", "dram_write_bytes", "source", "source_sha", "library_sha1_at_capture")}
    if sm_mhz:
        peak = sm_count * 4 * sm_mhz * 1e6          # 4 schedulers per SM, one warp instruction per clock each
        ach = fig["warp_instructions"] / (k_ms * 1e-3)
        roof["issue"] = {"bound": "issue slots (4 schedulers/SM)", "achieved": ach / 1e9, "peak": peak / 1e9, "unit": "G warp-instr/s",
                         "frac": ach / peak, "source_sha": fig.get("source_sha"), "library_sha1_at_capture": fig.get("library_sha1_at_capture"),
                         "note": "utilisation of the instructions the kernel executes, not a fraction of an algorithmic minimum; the "
                                 "instruction count is ncu's for a build of the sources with this hash (= the loaded library's sources)"}


def workload_name(a):
    if a.profile == "illumina":
        return ("synthetic Illumina-profile %d bp reads (per base: 2.452%% mismatch, 0.0468%% insert, 0.0553%% delete), "
                "%d pairs/batch/GPU, greedy k=%d x=o=e=1" % (a.length, a.pairs, a.k))
    return "simulated %d bp reads, error rate %.2f (exact edits, 96%% mismatch), %d pairs/batch/GPU, greedy k=%d x=o=e=1" % (
        a.length, a.error, a.pairs, a.k)


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 100 ms while the GPU legs run."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index = index
        self.proc = None
        self.lines = []
        self.on = False      # samples are kept only while a GPU leg (warm-up + timed region) is running

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._pump, daemon=True).start()
        except OSError:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.lines.append((self.on, line.strip()))

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.25)
        self.proc.terminate()
        sm, sm_load, mx, reasons = [], [], [], set()
        for on, ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1])); mx.append(float(f[2]))
                if on:
                    sm_load.append(float(f[1]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        # the timed regions are ~10 ms long and nvidia-smi samples every 100 ms: samples that fell inside a GPU leg
        # (generation + warm-up + timed region) are counted apart; the median is theirs when there are at least 3
        use = sm_load if len(sm_load) >= 3 else sm
        return {"sm_mhz": float(np.median(use)) if use else None, "sm_max_mhz": max(mx) if mx else None,
                "samples": len(sm), "samples_in_gpu_legs": len(sm_load), "sm_mhz_min": min(sm) if sm else None,
                "reasons": sorted(reasons)}


def load_pkg():
    import __graft_entry__ as ge
    return ge.load()


def measured_peak():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


# ------------------------------------------------------------------------------------------------
# CPU side.  This is the one place bench.py may execute anything under oracle/:
#   * the CHECKER (parity of the timed outputs, outside every timed region): the plain-C port oracle/gasma_oracle.c,
#     which is pinned against the unmodified reference by tests/test_oracle.py;
#   * the CPU BASELINE: the unmodified reference (oracle/_ref, kind "reference") when the workload is inside its
#     128-character domain, else the fast CPU arm oracle/gasma_fast.cpp (kind "port": W fixed at compile time, 64-bit
#     words, pext/popcnt; bit-equal to the checker and no slower than the reference at W=2 -- NOT the slow checker).
# ------------------------------------------------------------------------------------------------
def _ol():
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import oracle_lib as ol
    return ol


def host_threads():
    return len(os.sched_getaffinity(0))


def parity_greedy(batch_np, k, W, got, label, threads=None):
    """Full comparison of one batch's greedy output (cost, op_off, ops as numpy arrays) with the checker."""
    ol = _ol()
    b = ol.Batch(*batch_np)
    t0 = time.perf_counter()
    ref = ol.oracle_greedy(b, ol.make_params(k, 1, 1, 1, W, 1), threads=threads or host_threads())
    n_ops = int(ref["op_off"][-1])
    got_ops = got["ops"][:int(got["op_off"][b.n])]
    res = {"what": label, "pairs": int(b.n),
           "cost_mismatch": int((got["cost"][:b.n] != ref["cost"]).sum()),
           "op_off_mismatch": int((got["op_off"][:b.n + 1] != ref["op_off"]).sum()),
           "ops_mismatch": int(n_ops != len(got_ops)) if n_ops != len(got_ops) else int((got_ops != ref["ops"]).sum()),
           "oracle": "oracle/gasma_oracle.c (plain-C port, W=%d; pinned against the unmodified reference at W=2)" % W,
           "checker_seconds": round(time.perf_counter() - t0, 2)}
    res["ok"] = res["cost_mismatch"] == 0 and res["op_off_mismatch"] == 0 and res["ops_mismatch"] == 0
    return res


def cpu_greedy_baseline(batch_np, k, max_length, target_seconds=10.0, threads=None):
    ol = _ol()
    threads = threads or host_threads()
    b = ol.Batch(*batch_np)
    if max_length <= 128 and ol.ref_available():
        def run(bb):
            ol.ref_greedy(bb, k, 1, 1, 1, mode=1, threads=threads, want_text=False)
        kind = "reference"
        what = "unmodified hurdle_matrix<int_128bit> reset+run+get_cost+get_CIGAR (oracle/_ref, benchmark mode)"
    elif max_length <= 256:
        ol.build_oracle()

        def run(bb):
            ol.fast_greedy(bb, k, 1, 1, 1, max_length // 64, 1, threads=threads, want_ops=False)
        kind = "port"
        what = ("fast CPU arm oracle/gasma_fast.cpp, W=%d fixed at compile time (the reference itself stops at 128 "
                "characters); bit-equal to the checker, 1.7x the unmodified reference per thread at W=2" % (max_length // 64))
    else:
        ol.build_oracle()
        params = ol.make_params(k, 1, 1, 1, max_length // 64, 1)
        import ctypes as C

        def run(bb):
            cost = np.zeros(bb.n, np.int32)
            off = np.zeros(bb.n + 1, np.uint32)
            rc = ol.oracle().go_greedy_batch(ol._ptr(bb.reads), ol._ptr(bb.read_off), ol._ptr(bb.refs), ol._ptr(bb.ref_off),
                                             bb.n, C.byref(params), threads, ol._ptr(cost), None, 0, None, ol._ptr(off), 0, None)
            assert rc == 0
        kind = "port"
        what = "plain-C W=%d checker port (oracle/gasma_oracle.c): no fast CPU arm exists beyond 256 characters" % (max_length // 64)
    return _timed_sample(b, run, target_seconds, threads, kind, what)


def cpu_nw_baseline(batch_np, target_seconds=8.0, threads=None):
    """The verifier on the CPU (benchmark_utils.h:130-150): scalar Gotoh restatement, penalty + traceback, T threads.
    parasail itself is an absent submodule of the reference (README: 27-32 k pairs/s/core with its SSE kernel)."""
    ol = _ol()
    threads = threads or host_threads()
    b = ol.Batch(*batch_np)

    def run(bb):
        ol.oracle_nw(bb, 1, 1, 1, threads=threads, want_ops=True)
    return _timed_sample(b, run, target_seconds, threads, "port",
                         "scalar Gotoh NW + traceback (oracle/gasma_oracle.c:go_nw); parasail is not in the reference tree")


def _timed_sample(b, run, target_seconds, threads, kind, what):
    """Whole passes over a bounded sample of the batch until `target_seconds` of CPU work are done."""
    small = b.slice(0, min(b.n, max(64, 2000 * threads)))
    run(small)                                   # warm-up: thread start, page faults, library load
    t0 = time.perf_counter(); run(small); t1 = time.perf_counter() - t0
    rate = small.n / max(t1, 1e-9)
    sample = b.slice(0, int(min(b.n, max(small.n, rate * target_seconds / 4))))
    reps, t0 = 0, time.perf_counter()
    while True:
        run(sample)
        reps += 1
        dt = time.perf_counter() - t0
        if dt >= target_seconds or reps >= 1000:
            break
    return {"value": sample.n * reps / dt, "unit": "alignments/s", "cores": threads, "kind": kind,
            "sample": "%d pairs x %d passes of the same synthetic workload, %.1f s wall, %s" % (sample.n, reps, dt, what)}


def device_batch_to_numpy(t, n=None):
    """Host copy of a device batch (or of its first n pairs)."""
    d_reads, d_roff, d_refs, d_foff = t[:4]
    roff = d_roff.cpu().numpy().view(np.uint32)
    foff = d_foff.cpu().numpy().view(np.uint32)
    if n is not None:
        roff, foff = roff[:n + 1], foff[:n + 1]
    return (d_reads[:int(roff[-1])].cpu().numpy(), roff, d_refs[:int(foff[-1])].cpu().numpy(), foff)


# ------------------------------------------------------------------------------------------------
# One greedy leg: W warm-up + K timed steps of gasma_greedy_batch_dev over batches resident in HBM (cycled, each
# larger than L2 or flushed by its neighbour), CUDA events on the launching stream, max over ranks.  Then, outside
# the timed region: parity of the LAST step's output against the checker (whole batch or a prefix), the roofline
# block of the greedy kernel, and (1 GPU only) the CPU baseline on a bounded sample of the same batch.
# ------------------------------------------------------------------------------------------------
class Env:
    def __init__(self, torch, dist, pkg, dev, local, rank, world, stream, sampler=None):
        self.torch, self.dist, self.pkg = torch, dist, pkg
        self.dev, self.local, self.rank, self.world, self.stream = dev, local, rank, world, stream
        self.sampler = sampler

    def load(self, on):
        """Clock samples count only while a GPU leg is under load (not during the CPU checker / baseline phases)."""
        if self.sampler is not None:
            if not on:
                time.sleep(0.05)
            self.sampler.on = on

    def barrier(self):
        if self.world > 1:
            self.dist.barrier()
        self.torch.cuda.synchronize()

    def max_over_ranks(self, v):
        t = self.torch.tensor([v], device=self.dev, dtype=self.torch.float64)
        if self.world > 1:
            self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return float(t.item())

    def sum_over_ranks(self, vals):
        t = self.torch.tensor(vals, device=self.dev, dtype=self.torch.int64)
        if self.world > 1:
            self.dist.all_reduce(t, op=self.dist.ReduceOp.SUM)   # the path's only collective: a handful of counters
        return [int(v) for v in t.tolist()]


def greedy_leg(env, *, workload, length, error, profile, k, pairs, steps, warmup, seed, n_batches=2,
               parity_pairs=None, cpu_seconds=0.0, want_parity=True, in_flight=0):
    torch, pkg = env.torch, env.pkg
    ml = max_length_for(length, error if profile == "simulated" else 0.0)
    al = pkg.Aligner(k=k, max_length=ml, device=env.local)
    al.set_stream(env.stream.cuda_stream)
    if profile == "illumina":
        cfg = pkg.gen_config(seed + 16, length, 0.0, mode=2, profile=PROFILE_ILLUMINA)
    else:
        cfg = pkg.gen_config(seed, length, error)
    n, dev = pairs, env.dev
    env.load(True)      # generation, warm-up and the timed region all keep the GPU busy
    batches = []
    for bi in range(n_batches):
        d_reads = torch.empty(n * length + 16, dtype=torch.uint8, device=dev)
        d_refs = torch.empty(al.generate_ref_capacity(cfg, n), dtype=torch.uint8, device=dev)
        d_roff = torch.empty(n + 1, dtype=torch.int32, device=dev)
        d_foff = torch.empty(n + 1, dtype=torch.int32, device=dev)
        first = (env.rank * n_batches + bi) * n   # global pair index: shards never overlap
        al.generate_dev(cfg, first, n, d_reads, d_roff, d_refs, d_foff)
        batches.append((d_reads, d_roff, d_refs, d_foff))
    d_cost = torch.empty(n, dtype=torch.int32, device=dev)
    ops_per_pair = max(24, ml // 2 + 8)
    d_ops = torch.empty(n * ops_per_pair, dtype=torch.int16, device=dev)
    d_ooff = torch.empty(n + 1, dtype=torch.int32, device=dev)
    al.synchronize()

    def step(i):
        r, ro, f, fo = batches[i % len(batches)]
        al.greedy_batch_dev(r, ro, f, fo, n, d_cost, d_ops, d_ooff)

    for i in range(warmup):
        step(i)
    al.synchronize()
    al.kernel_time_ms(0)
    launches0 = al.launch_count
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    env.barrier()
    ev0.record(env.stream)
    for i in range(steps):
        step(i)
    ev1.record(env.stream)
    env.barrier()
    ms_rank = ev0.elapsed_time(ev1)
    ms_total = env.max_over_ranks(ms_rank)
    al.synchronize()
    env.load(False)
    launches = al.launch_count - launches0
    kernel_ms, kernel_launches = al.kernel_time_ms(0)
    k_ms = kernel_ms / max(kernel_launches, 1)
    if os.environ.get("GASMA_BENCH_RANKS"):   # per-rank view of the leg on stderr (scaling diagnostics)
        sys.stderr.write("[rank %d] %s: step %.4f ms, greedy kernel %.4f ms\n" % (env.rank, workload[:40], ms_rank / steps, k_ms))

    last = batches[(steps - 1) % len(batches)]
    ooff_last = int(d_ooff[-1].item()) & 0xFFFFFFFF
    rbytes, fbytes = int(last[1][-1].item()), int(last[3][-1].item())
    # algorithmic bytes of one batch: m + n + 4 (cost) + 4 (CSR offset) + CIGAR bytes produced
    alg_bytes = rbytes + fbytes + 8 * n + 2 * ooff_last
    tot = env.sum_over_ranks([n, int(d_cost.sum(dtype=torch.int64).item()), ooff_last])
    # ---- the same K steps with several batches in flight: one context (own scratch, own stream, own output buffers)
    # per batch in flight, round-robin.  A greedy launch ends with a drain (a thread works on ~9 pairs per launch, the
    # last one finishes alone) followed by the compaction kernel and the next launch's ramp-up; with the next batch
    # already queued on another stream those gaps are filled.  `value` stays the one-batch-at-a-time number, which the
    # kernel time and the roofline block belong to. ----
    flight = None
    if in_flight > 1:
        streams = [env.stream] + [torch.cuda.Stream(device=dev) for _ in range(in_flight - 1)]
        als = [al]
        for s_ in streams[1:]:
            x_ = pkg.Aligner(k=k, max_length=ml, device=env.local)
            x_.set_stream(s_.cuda_stream)
            als.append(x_)
        outs = [(torch.empty_like(d_cost), torch.empty_like(d_ops), torch.empty_like(d_ooff)) for _ in als]

        def fstep(i):
            r, ro, f, fo = batches[i % len(batches)]
            c_ = i % len(als)
            als[c_].greedy_batch_dev(r, ro, f, fo, n, *outs[c_])

        env.load(True)
        for i in range(max(warmup, 2 * len(als))):
            fstep(i)
        torch.cuda.synchronize()
        env.barrier()
        f0, f1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        f0.record(env.stream)
        for s_ in streams[1:]:
            s_.wait_event(f0)
        for i in range(steps):
            fstep(i)
        for s_ in streams[1:]:
            e_ = torch.cuda.Event()
            e_.record(s_)
            env.stream.wait_event(e_)
        f1.record(env.stream)
        env.barrier()
        f_ms = env.max_over_ranks(f0.elapsed_time(f1))
        env.load(False)
        for x_ in als:
            x_.synchronize()
        c_ = (steps - 1) % len(als)       # the last step ran the batch the sequential pass ended on: same outputs?
        same = bool(torch.equal(outs[c_][0], d_cost) and torch.equal(outs[c_][2], d_ooff) and
                    torch.equal(outs[c_][1][:ooff_last], d_ops[:ooff_last]))
        flight = {"batches_in_flight": len(als), "value": env.world * n * steps / (f_ms * 1e-3), "unit": "alignments/s",
                  "ms_per_step": f_ms / steps, "steps": steps, "outputs_equal_sequential": same,
                  "note": "the same K steps handed round-robin to %d contexts, each on its own stream with its own scratch and "
                          "output buffers (tools/two_stream_probe.py); kernel times overlap, so no roofline block belongs to it" % len(als)}
        for x_ in als[1:]:
            x_.close()
        del outs
    peak, peak_src = measured_peak()
    achieved = alg_bytes / (k_ms * 1e-3) / 1e9
    W = ml // 64
    shape = "greedy_kernel<W=%d> (one alignment per thread, persistent warps)" % W if ml <= 256 else \
            "greedy_warp_kernel<W=%d> (one alignment per warp)" % W
    res = {"workload": workload, "max_length": ml, "value": env.world * n * steps / (ms_total * 1e-3), "unit": "alignments/s",
           "steps": steps, "warmup": warmup, "ms_per_step": ms_total / steps, "gpu_launches": int(launches),
           "cache": "inputs larger than L2 (%.0f MB per batch, %d batches cycled)" % ((rbytes + fbytes) / 1e6, n_batches)
                    if (rbytes + fbytes) * n_batches > 252e6 else
                    "%d batches of %.0f MB cycled: together smaller than 2 x L2, treat as L2-warm" % (n_batches, (rbytes + fbytes) / 1e6),
           "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": None,
                        "kernel": shape, "kernel_ms": k_ms, "kernel_share_of_step": k_ms / (ms_total / steps),
                        "algorithmic_bytes_per_launch": alg_bytes, "peak_source": peak_src},
           "counters": {"pairs": tot[0], "penalty_sum": tot[1], "cigar_ops": tot[2]}}
    if flight is not None:
        res["in_flight"] = flight
    state = {"al": al, "batches": batches, "last": last, "d_cost": d_cost, "d_ops": d_ops, "d_ooff": d_ooff, "ml": ml,
             "ops_per_pair": ops_per_pair, "n": n, "k": k, "alg_bytes": alg_bytes, "k_ms": k_ms}
    if env.rank == 0 and want_parity:
        np_n = n if parity_pairs is None else min(n, parity_pairs)
        hb = device_batch_to_numpy(last, np_n)
        got = {"cost": d_cost[:np_n].cpu().numpy(), "op_off": d_ooff[:np_n + 1].cpu().numpy().view(np.uint32)}
        got["ops"] = d_ops[:int(got["op_off"][np_n])].cpu().numpy().view(np.uint16)
        res["parity"] = parity_greedy(hb, k, W, got, "device-resident output of the last timed step, pairs [0, %d) of the batch" % np_n)
        if flight is not None:
            res["parity"]["in_flight_outputs_equal_sequential"] = flight["outputs_equal_sequential"]
            res["parity"]["ok"] = bool(res["parity"]["ok"] and flight["outputs_equal_sequential"])
    if env.rank == 0 and env.world == 1 and cpu_seconds > 0:
        res["cpu_baseline"] = cpu_greedy_baseline(device_batch_to_numpy(last, min(n, 400_000)), k, ml, target_seconds=cpu_seconds)
    return res, state


def pick_call(res):
    """`value` of an e2e block = the faster of the two public ways to make the call (one synchronous call per step, or
    the asynchronous pair with two calls in flight); both numbers stay in the block."""
    ap = res["async_pair"]
    if ap["equals_sync_call"] and ap["value"] > res["value"]:
        res["value"] = ap["value"]
        res["h2d_gb_per_s"] *= ap["value"] / res["sync_call_value"]
        res["call"] = "asynchronous pair (gasma_greedy_submit* / gasma_greedy_wait), two calls in flight"
    else:
        res["call"] = "one synchronous call per step"


def e2e_leg(env, st, steps, warmup, want_parity=True):
    """The host-pointer C-ABI call gasma_greedy_batch on pinned host buffers: H2D of both CSR strings, all kernels,
    D2H of penalties + CSR CIGAR inside the timed region, every step."""
    torch = env.torch
    al, n = st["al"], st["n"]
    hb = device_batch_to_numpy(st["batches"][0])
    pin = lambda arr: torch.from_numpy(np.ascontiguousarray(arr)).pin_memory()
    h_reads, h_roff, h_refs, h_foff = [pin(x.view(np.uint8) if x.dtype == np.uint8 else x.view(np.int32)) for x in hb]
    h_cost = torch.empty(n, dtype=torch.int32).pin_memory()
    h_ops = torch.empty(n * st["ops_per_pair"], dtype=torch.int16).pin_memory()
    h_ooff = torch.empty(n + 1, dtype=torch.int32).pin_memory()
    out = {"cost": h_cost.numpy(), "ops": h_ops.numpy().view(np.uint16), "op_off": h_ooff.numpy().view(np.uint32)}
    np_in = (h_reads.numpy(), h_roff.numpy().view(np.uint32), h_refs.numpy(), h_foff.numpy().view(np.uint32))

    def e2e_step():
        al.greedy_batch(np_in[0], np_in[1], np_in[2], np_in[3], want_ops=True, ops_cap=len(out["ops"]), out=out)

    env.load(True)
    for _ in range(max(1, min(warmup, 3))):
        e2e_step()
    e2e_steps = max(3, min(steps, 10))
    env.barrier()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        e2e_step()
    torch.cuda.synchronize()
    e2e_s = env.max_over_ranks(time.perf_counter() - t0)
    env.load(False)
    n_ops_host = int(out["op_off"][n])
    h2d = int(np_in[0].nbytes + np_in[2].nbytes + np_in[1].nbytes + np_in[3].nbytes)
    d2h = int(4 * n + 4 * (n + 1) + 2 * n_ops_host)
    res = {"value": env.world * n * e2e_steps / e2e_s, "unit": "alignments/s", "h2d_bytes_per_step": h2d,
           "d2h_bytes_per_step": d2h, "steps": e2e_steps, "h2d_gb_per_s": h2d * e2e_steps / e2e_s / 1e9,
           "note": "host-pointer C-ABI call, pinned buffers, 8 pipelined slices; bound by the host-to-device link "
                   "(ASCII in; PCIe 5 x16 measured at 55 GB/s by tools/link_probe.py)"}
    def pinned_out():
        ts = [torch.empty_like(h_cost).pin_memory(), torch.empty_like(h_ops).pin_memory(), torch.empty_like(h_ooff).pin_memory()]
        return {"cost": ts[0].numpy(), "ops": ts[1].numpy().view(np.uint16), "op_off": ts[2].numpy().view(np.uint32), "_pin": ts}

    def same_as_sync(o):
        return bool(np.array_equal(o["cost"], out["cost"]) and np.array_equal(o["op_off"], out["op_off"]) and
                    np.array_equal(o["ops"][:n_ops_host], out["ops"][:n_ops_host]))

    def async_pair_leg(submit, what):
        """The same call as an asynchronous pair (gasma_greedy_submit / gasma_greedy_wait): two calls in flight, each
        with its own pinned result buffers; every step still copies its inputs in and its results out."""
        outs = [pinned_out(), pinned_out()]

        def loop(k):
            pend = []
            for i in range(k):
                if len(pend) == 2:
                    al.greedy_wait(pend.pop(0))
                pend.append(submit(outs[i & 1]))
            for t in pend:
                al.greedy_wait(t)

        env.load(True)
        loop(max(2, min(warmup, 4)))
        env.barrier()
        t0 = time.perf_counter()
        loop(e2e_steps)
        torch.cuda.synchronize()
        s = env.max_over_ranks(time.perf_counter() - t0)
        env.load(False)
        return {"value": env.world * n * e2e_steps / s, "unit": "alignments/s", "steps": e2e_steps, "calls_in_flight": 2,
                "equals_sync_call": same_as_sync(outs[0]) and same_as_sync(outs[1]), "what": what}

    res["sync_call_value"] = res["value"]
    res["async_pair"] = async_pair_leg(
        lambda o: al.greedy_submit(np_in[0], np_in[1], np_in[2], np_in[3], want_ops=True, out=o),
        "gasma_greedy_submit / gasma_greedy_wait on the same pinned ASCII buffers")
    pick_call(res)
    parity = None
    if env.rank == 0 and want_parity:
        parity = parity_greedy(hb, st["k"], st["ml"] // 64, out, "host buffers written by the last timed e2e call (whole batch)")
        parity["async_pair_equals_sync_call"] = res["async_pair"]["equals_sync_call"]
        parity["ok"] = bool(parity["ok"] and res["async_pair"]["equals_sync_call"])
    # ---- the same batch, pre-packed OUTSIDE the timed region (2-bit streams; the reference's batch driver
    # LEAP_SIMD/main.cpp:108-140 converts before its timed loop too): H2D of the four streams + offsets, kernels, D2H ----
    packed = None
    if st["ml"] <= 256:
        r0, r1 = al.pack2bit_stream(np_in[0])
        f0, f1 = al.pack2bit_stream(np_in[2])
        pk = [pin(x.view(np.int16)) for x in (r0, r1, f0, f1)]
        pk_np = [t.numpy().view(np.uint16) for t in pk]
        out2 = {"cost": np.empty_like(out["cost"]), "ops": np.empty_like(out["ops"]), "op_off": np.empty_like(out["op_off"])}
        h2 = [torch.from_numpy(out2[k].view(np.int16) if out2[k].dtype == np.uint16 else out2[k].view(np.int32)).pin_memory() for k in ("cost", "ops", "op_off")]
        out2 = {"cost": h2[0].numpy(), "ops": h2[1].numpy().view(np.uint16), "op_off": h2[2].numpy().view(np.uint32)}

        def packed_step():
            al.greedy_batch_packed(pk_np[0], pk_np[1], np_in[1], pk_np[2], pk_np[3], np_in[3], want_ops=True, out=out2)

        env.load(True)
        for _ in range(max(1, min(warmup, 3))):
            packed_step()
        env.barrier()
        t0 = time.perf_counter()
        for _ in range(e2e_steps):
            packed_step()
        torch.cuda.synchronize()
        p_s = env.max_over_ranks(time.perf_counter() - t0)
        env.load(False)
        h2d_p = int(sum(x.nbytes for x in pk_np) + np_in[1].nbytes + np_in[3].nbytes)
        same = bool(np.array_equal(out2["cost"], out["cost"]) and np.array_equal(out2["op_off"], out["op_off"]) and
                    np.array_equal(out2["ops"][:n_ops_host], out["ops"][:n_ops_host]))
        packed = {"value": env.world * n * e2e_steps / p_s, "unit": "alignments/s", "h2d_bytes_per_step": h2d_p, "d2h_bytes_per_step": d2h,
                  "steps": e2e_steps, "h2d_gb_per_s": h2d_p * e2e_steps / p_s / 1e9, "equals_ascii_call": same,
                  "note": "gasma_greedy_batch_packed on pinned host buffers: inputs as 2-bit streams packed outside the timed region "
                          "(%.0f B per pair instead of %.0f B of ASCII)" % (h2d_p / n, h2d / n)}
        packed["sync_call_value"] = packed["value"]
        packed["async_pair"] = async_pair_leg(
            lambda o: al.greedy_submit_packed(pk_np[0], pk_np[1], np_in[1], pk_np[2], pk_np[3], np_in[3], want_ops=True, out=o),
            "gasma_greedy_submit_packed / gasma_greedy_wait on the same pinned 2-bit streams")
        pick_call(packed)
        if parity is not None:
            parity["packed_async_pair_equals_sync_call"] = packed["async_pair"]["equals_sync_call"]
            parity["ok"] = bool(parity["ok"] and packed["async_pair"]["equals_sync_call"] and same)
    res["packed"] = packed
    return res, parity


# ------------------------------------------------------------------------------------------------
# Secondary leg: the whole benchmark body (BASELINE.json configs[2]: error rate 0.20, greedy + NW verifier +
# accuracy / coverage counters), batches resident in HBM, counters summed over ranks with one all-reduce.
# ------------------------------------------------------------------------------------------------
def full_path_leg(env, a, L=100, e=0.20, seed_off=2, label="configs[2]"):
    torch, pkg, dev = env.torch, env.pkg, env.dev
    from gasma_b200 import shard
    n, steps = a.pairs, max(3, min(a.steps, 6))
    al = pkg.Aligner(k=a.k, max_length=128, device=env.local)
    al.set_stream(env.stream.cuda_stream)
    cfg = pkg.gen_config(SEED + seed_off, L, e)
    counters = torch.zeros(4, dtype=torch.int64, device=dev)
    # every rank takes its own batches of the run (round-robin by batch index, keyed by the global pair index);
    # all of them are generated up front so that the timed steps see inputs resident in HBM (~210 MB per batch > L2)
    sched = shard.batch_schedule((steps + 1) * n * env.world, n, env.rank, env.world)
    batches = []
    for first, cnt in sched:
        d_reads = torch.empty(n * L + 16, dtype=torch.uint8, device=dev)
        d_refs = torch.empty(al.generate_ref_capacity(cfg, n), dtype=torch.uint8, device=dev)
        d_roff = torch.empty(n + 1, dtype=torch.int32, device=dev)
        d_foff = torch.empty(n + 1, dtype=torch.int32, device=dev)
        al.generate_dev(cfg, first, cnt, d_reads, d_roff, d_refs, d_foff)
        batches.append((d_reads, d_roff, d_refs, d_foff, cnt))
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    env.load(True)
    for i, (d_reads, d_roff, d_refs, d_foff, cnt) in enumerate(batches):
        if i == 1:      # batch 0 is the warm-up (allocations, first launches)
            counters.zero_()
            al.synchronize(); al.kernel_time_ms(0); al.kernel_time_ms(1)
            env.barrier()
            ev0.record(env.stream)
        al.benchmark_batch_dev(d_reads, d_roff, d_refs, d_foff, cnt, counters)
    ev1.record(env.stream)
    torch.cuda.synchronize()
    al.synchronize()
    ms = env.max_over_ranks(ev0.elapsed_time(ev1))
    g_ms, g_n = al.kernel_time_ms(0)
    nw_ms, nw_n = al.kernel_time_ms(1)
    # ---- the same batches with several of them in flight: one context (own scratch) and one stream per batch in
    # flight, round-robin.  The verifier of one batch then shares the GPU with the greedy kernel of the next and the
    # tails of both grids are filled (tools/two_stream_probe.py).  Same kernels, same counters. ----
    flight = None
    if a.in_flight > 1:
        streams = [env.stream] + [torch.cuda.Stream(device=dev) for _ in range(a.in_flight - 1)]
        als = [al]
        for s_ in streams[1:]:
            x_ = pkg.Aligner(k=a.k, max_length=128, device=env.local)
            x_.set_stream(s_.cuda_stream)
            als.append(x_)
        cnts = [torch.zeros(4, dtype=torch.int64, device=dev) for _ in als]
        for i, x_ in enumerate(als):      # warm-up of the new contexts (allocations), on the warm-up batch
            x_.benchmark_batch_dev(*batches[0][:4], batches[0][4], cnts[i])
        torch.cuda.synchronize()
        for c_ in cnts:
            c_.zero_()
        torch.cuda.synchronize()
        env.barrier()
        f0, f1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        f0.record(env.stream)
        for s_ in streams[1:]:
            s_.wait_event(f0)
        for i, (d_reads, d_roff, d_refs, d_foff, cnt) in enumerate(batches[1:]):
            als[i % len(als)].benchmark_batch_dev(d_reads, d_roff, d_refs, d_foff, cnt, cnts[i % len(als)])
        for s_ in streams[1:]:
            e_ = torch.cuda.Event()
            e_.record(s_)
            env.stream.wait_event(e_)
        f1.record(env.stream)
        torch.cuda.synchronize()
        for x_ in als:
            x_.synchronize()
        f_ms = env.max_over_ranks(f0.elapsed_time(f1))
        f_cnt = [int(v) for v in sum(cnts).tolist()]
        flight = {"batches_in_flight": len(als), "ms_per_step": f_ms / steps,
                  "counters_equal_sequential": f_cnt == [int(v) for v in counters.tolist()]}
        for x_ in als[1:]:
            x_.close()
    # the report's LEAP column (comparator): its own kernel over the same resident batches, timed apart
    d_nw = torch.empty(n, dtype=torch.int32, device=dev)
    d_gc = torch.empty(n, dtype=torch.int32, device=dev)
    d_leap = torch.empty(n, dtype=torch.int32, device=dev)
    scratch = torch.zeros(4, dtype=torch.int64, device=dev)
    leap_ok = torch.zeros(1, dtype=torch.int64, device=dev)
    al.kernel_time_ms(2)
    for (d_reads, d_roff, d_refs, d_foff, cnt) in batches[1:]:
        al.benchmark_batch_dev(d_reads, d_roff, d_refs, d_foff, cnt, scratch, d_nw_penalty=d_nw)
        al.leap_batch_dev(d_reads, d_roff, d_refs, d_foff, cnt, d_leap)
        al.synchronize()
        leap_ok += (d_leap[:cnt] == d_nw[:cnt]).sum()
    leap_ms, leap_n = al.kernel_time_ms(2)
    env.load(False)
    c = env.sum_over_ranks([int(v) for v in counters.tolist()] + [int(leap_ok.item())])
    last = batches[-1]
    rb, fb = int(last[1][-1].item()), int(last[3][-1].item())
    peak, peak_src = measured_peak()
    nw_k = nw_ms / max(nw_n, 1)
    nw_alg = rb + fb + 12 * n        # SURVEY.md 8d for the verifier: m + n read once, greedy penalty in, NW penalty out, counters
    res = {"workload": "simulated %d bp reads, error rate %.2f, %d pairs/batch/GPU, greedy + NW verifier + accuracy/coverage counters, "
                       "batches resident in HBM (%s)" % (L, e, n, label),
           "value": c[0] / (ms * 1e-3), "unit": "alignments/s", "steps": steps, "ms_per_step": ms / steps,
           "batches_in_flight": 1, "greedy_kernel_ms": g_ms / max(g_n, 1), "nw_kernel_ms": nw_k,
           "roofline": {"bound": "hbm", "achieved": nw_alg / (nw_k * 1e-3) / 1e9, "peak": peak, "unit": "GB/s",
                        "frac": nw_alg / (nw_k * 1e-3) / 1e9 / peak, "traffic": None, "kernel": "nw_band_kernel<W=2,64,counters> (+ deferred / hard-cover tails)",
                        "kernel_ms": nw_k, "kernel_share_of_step": nw_k / (ms / steps), "algorithmic_bytes_per_launch": nw_alg,
                        "peak_source": peak_src, "note": "bit-vector DP: bound by instruction issue, not HBM"},
           "counters": {"total": c[0], "nw_correct": c[1], "greedy_correct": c[2], "greedy_coverage": c[3], "leap_correct": c[4]},
           "accuracy_pct": 100.0 * c[2] / max(c[0], 1), "coverage_pct": 100.0 * c[3] / max(c[0], 1),
           "leap": {"kernel_ms": leap_ms / max(leap_n, 1), "accuracy_pct": 100.0 * c[4] / max(c[0], 1),
                    "readme_accuracy_pct": 88.579, "note": "comparator column of the report, timed outside the step"},
           "readme_reference": {"accuracy_pct": 46.023, "coverage_pct": 88.289, "source": "README.md:65,67 (e=0.20, other random data)"}}
    if flight is not None:
        # `value` = the faster of the two ways to drive the same call; the one-batch-at-a-time numbers (which the kernel
        # times and the roofline block belong to) stay beside it
        res["sequential"] = {"value": res["value"], "ms_per_step": res["ms_per_step"]}
        flight["value"] = c[0] / (flight["ms_per_step"] * steps * 1e-3)
        res["in_flight"] = flight
        if flight["counters_equal_sequential"] and flight["value"] > res["value"]:
            res["value"], res["ms_per_step"], res["batches_in_flight"] = flight["value"], flight["ms_per_step"], flight["batches_in_flight"]
            res["roofline"]["kernel_share_of_step"] = None
            res["roofline"]["note"] += "; kernel_ms is the verifier's duration with one batch in flight"
    if env.rank == 0 and not a.no_parity:
        # parity of the whole body on a prefix of the last batch: the same call on the first S pairs, against the checker's
        # greedy + Gotoh NW + metrics + LEAP on the same pairs (NW with traceback is ~30 k pairs/s/core: bounded sample)
        S = min(n, 100_000)
        ol = _ol()
        cnt4 = torch.zeros(4, dtype=torch.int64, device=dev)
        al.benchmark_batch_dev(last[0], last[1], last[2], last[3], S, cnt4, d_greedy_cost=d_gc, d_nw_penalty=d_nw)
        al.leap_batch_dev(last[0], last[1], last[2], last[3], S, d_leap)
        al.synchronize()
        hb = ol.Batch(*device_batch_to_numpy(last, S))
        t0 = time.perf_counter()
        T = host_threads()
        og = ol.oracle_greedy(hb, ol.make_params(a.k, 1, 1, 1, 2, 1), threads=T)
        on = ol.oracle_nw(hb, 1, 1, 1, threads=T)
        om = ol.oracle_metrics(hb, og, on)
        olp = ol.oracle_leap(hb, a.k, 1, 1, 1, threads=T)
        got = [int(v) for v in cnt4.tolist()]
        res["parity"] = {"what": "gasma_benchmark_batch_dev + gasma_leap_batch_dev on pairs [0, %d) of the last timed batch" % S,
                         "pairs": S, "counters_gpu": got, "counters_oracle": [int(v) for v in om.tolist()],
                         "greedy_cost_mismatch": int((d_gc[:S].cpu().numpy() != og["cost"]).sum()),
                         "nw_penalty_mismatch": int((d_nw[:S].cpu().numpy() != on["penalty"]).sum()),
                         "leap_mismatch": int((d_leap[:S].cpu().numpy() != olp).sum()),
                         "oracle": "oracle/gasma_oracle.c: go_greedy + go_nw (scalar Gotoh, traceback rule diagonal > I > D) + go_metrics + go_leap",
                         "checker_seconds": round(time.perf_counter() - t0, 2)}
        p = res["parity"]
        p["ok"] = (p["counters_gpu"] == p["counters_oracle"] and p["greedy_cost_mismatch"] == 0 and
                   p["nw_penalty_mismatch"] == 0 and p["leap_mismatch"] == 0 and
                   (flight is None or flight["counters_equal_sequential"]))
        if env.world == 1 and not a.no_cpu_baseline:
            res["cpu_baseline"] = {"nw_verifier": cpu_nw_baseline(device_batch_to_numpy(last, 200_000)),
                                   "greedy": cpu_greedy_baseline(device_batch_to_numpy(last, 400_000), a.k, 128, target_seconds=5.0)}
    al.close()
    return res


# ------------------------------------------------------------------------------------------------
# B200 arm
# ------------------------------------------------------------------------------------------------
def run_b200(a):
    import torch
    import torch.distributed as dist
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py --impl b200 needs a B200: there is no CPU fallback")
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    pkg = load_pkg()
    from gasma_b200 import shard
    numa = None if os.environ.get("GASMA_NO_NUMA_BIND") else shard.bind_host_to_device(local)
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    env = Env(torch, dist, pkg, torch.device("cuda", local), local, rank, world, torch.cuda.current_stream(),
              sampler if rank == 0 else None)
    want_cpu = not a.no_cpu_baseline and world == 1

    # ---- headline: BASELINE.json configs[1] (or whatever the flags name) ----
    head, st = greedy_leg(env, workload=workload_name(a), length=a.length, error=a.error, profile=a.profile, k=a.k, pairs=a.pairs,
                          steps=a.steps, warmup=a.warmup, seed=SEED + 1, n_batches=a.batches, parity_pairs=None,
                          cpu_seconds=10.0 if want_cpu else 0.0, want_parity=not a.no_parity, in_flight=a.in_flight)
    e2e, e2e_parity = e2e_leg(env, st, a.steps, a.warmup, want_parity=not a.no_parity)
    e2e["host_numa"] = {"node": numa[0], "cpus": numa[1]} if numa else None
    e2e_packed = e2e.pop("packed", None)
    ml = st["ml"]
    st["al"].close()
    del st
    torch.cuda.empty_cache()

    full = full_path_leg(env, a) if not a.no_full_path else None

    # ---- the other BASELINE.json configs, short legs under the same clock (value + roofline + parity + CPU sample) ----
    configs = {}
    if not a.no_configs:
        cs, cw = max(3, min(a.steps, 5)), max(3, min(a.warmup, 3))
        legs = [
            ("configs[0] README setup, in the reference's own 128-character domain", dict(length=100, error=0.05, profile="simulated", k=3, pairs=a.pairs), 200_000),
            ("configs[3] Illumina profile", dict(length=100, error=0.0, profile="illumina", k=3, pairs=a.pairs), 200_000),
            ("configs[4] long reads 1 kbp, e=0.15, k=24", dict(length=1000, error=0.15, profile="simulated", k=24, pairs=max(1000, a.pairs // 10)), 12_000),
            ("configs[4] long reads 1 kbp, e=0.15, k=50 (MAX_K)", dict(length=1000, error=0.15, profile="simulated", k=50, pairs=max(1000, a.pairs // 10)), 8_000),
        ]
        for i, (name, kw, ppairs) in enumerate(legs):
            aa = argparse.Namespace(**{**vars(a), **kw})
            r, s2 = greedy_leg(env, workload=name + ": " + workload_name(aa), steps=cs, warmup=cw, seed=SEED + 3 + i, n_batches=2,
                               parity_pairs=ppairs, cpu_seconds=4.0 if want_cpu else 0.0, want_parity=not a.no_parity, **kw)
            s2["al"].close()
            del s2
            torch.cuda.empty_cache()
            configs[name.split(":")[0]] = r
    clocks = sampler.stop() if rank == 0 else None   # samples of the GPU legs only (warm-up + timed regions, all under load)

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    roof = head["roofline"]
    sms = torch.cuda.get_device_properties(local).multi_processor_count
    attach_ncu(roof, ncu_figures(pkg, "headline"), roof["kernel_ms"], sms, clocks.get("sm_mhz") if clocks else None)
    if full is not None:
        attach_ncu(full["roofline"], ncu_figures(pkg, "full_path"), full["roofline"]["kernel_ms"], sms, clocks.get("sm_mhz") if clocks else None)
    roof["note"] = ("integer bit-vector work: the kernel is bound by instruction issue, not by HBM; frac is reported against the "
                    "HBM roofline as the contract asks (profiles/ holds the ncu issue / pipe figures of this build)")
    line = {
        "metric": "alignments/sec (greedy + banded NW) at 1/2/4/8 B200 vs AVX2 host cores",
        "value": head["value"], "unit": "alignments/s", "n_gpus": world, "steps": a.steps, "warmup": a.warmup,
        "ms_per_step": head["ms_per_step"], "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "u64 bit-vectors + int32 penalties (f64 only in the lane heuristic)", "data": "synthetic",
        "config": config_dict(a, ml), "cache": head["cache"], "e2e": e2e, "gpu_launches": head["gpu_launches"], "clocks": clocks,
        "roofline": roof, "counters": head["counters"], "library_sha1": lib_sha(pkg),
    }
    if "in_flight" in head:
        line["in_flight"] = head["in_flight"]
    line["source_sha"] = source_sha(pkg)
    if e2e_packed is not None:
        line["e2e_packed"] = e2e_packed
    if not a.no_parity:
        pr, pe = head.get("parity"), e2e_parity
        line["parity"] = {"ok": bool(pr and pe and pr["ok"] and pe["ok"] and (e2e_packed is None or e2e_packed["equals_ascii_call"])), "pairs": (pr["pairs"] if pr else 0) + (pe["pairs"] if pe else 0),
                          "cost_mismatch": (pr["cost_mismatch"] if pr else -1) + (pe["cost_mismatch"] if pe else -1),
                          "ops_mismatch": (pr["ops_mismatch"] + pr["op_off_mismatch"] if pr else -1) + (pe["ops_mismatch"] + pe["op_off_mismatch"] if pe else -1),
                          "oracle": pr["oracle"] if pr else None, "resident": pr, "e2e": pe,
                          "scope": "rank 0's shard (every rank runs the same kernels on its own pair-index range)"}
    if full is not None:
        line["full_path"] = full
    if configs:
        line["configs"] = configs
    if not a.no_parity and rank == 0:   # the verdict covers every leg that carries a parity block
        legs = ([full] if full is not None else []) + list(configs.values())
        line["parity"]["other_legs_ok"] = all(l["parity"]["ok"] for l in legs if "parity" in l)
        line["parity"]["ok"] = bool(line["parity"]["ok"] and line["parity"]["other_legs_ok"])
    if "cpu_baseline" in head:
        line["cpu_baseline"] = head["cpu_baseline"]
        c0 = configs.get("configs[0] README setup, in the reference's own 128-character domain")
        if c0 and "cpu_baseline" in c0:   # one ratio in the line against the reference's OWN code
            line["cpu_reference_L100"] = {"gpu_value": c0["value"], "cpu": c0["cpu_baseline"],
                                          "ratio_resident": c0["value"] / c0["cpu_baseline"]["value"]}
    if world > 1:
        dist.destroy_process_group()
    emit_line(line)
    if not a.no_parity and not line["parity"]["ok"]:
        raise SystemExit("PARITY FAILURE: the timed outputs differ from the oracle (see the parity block)")


# ------------------------------------------------------------------------------------------------
# reference arm: the reference's CPU implementation of the path on the box's host cores
# ------------------------------------------------------------------------------------------------
def run_reference(a):
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if rank != 0:
        return
    ol = _ol()
    ol.build_oracle()
    ml = max_length_for(a.length, a.error if a.profile == "simulated" else 0.0)
    threads = host_threads()
    # bounded sample of the same workload, regenerated from the same (seed, pair index) counters by the CPU twin
    n = min(a.pairs, 200_000)
    cfg = gen_config_for(ol, a, SEED + 1)
    b = ol.generate(cfg, 0, n)
    batch = (b.reads, b.read_off, b.refs, b.ref_off)
    per_step = []
    base = None
    for i in range(a.warmup + a.steps):
        base = cpu_greedy_baseline(batch, a.k, ml, target_seconds=max(1.0, min(6.0, 120.0 / max(1, a.warmup + a.steps))), threads=threads)
        if i >= a.warmup:
            per_step.append(base["value"])
    v = float(np.mean(per_step))
    base["value"] = v
    line = {"impl": "reference", "metric": "alignments/sec (greedy + banded NW) at 1/2/4/8 B200 vs AVX2 host cores",
            "value": v, "unit": "alignments/s", "n_gpus": world, "steps": a.steps, "warmup": a.warmup,
            "ms_per_step": 1e3 * n / v, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "u64/SSE bit-vectors + int32 penalties", "data": "synthetic",
            "config": config_dict(a, ml), "cpu_baseline": base,
            "e2e": {"value": v, "unit": "alignments/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    emit_line(line)


_REAL_STDOUT = None


def emit_line(line):
    """The ONE JSON line goes to the process's real stdout; everything libraries print on fd 1 while the run lasts
    (NCCL's version banner under torchrun) was sent to stderr by __main__."""
    data = (json.dumps(line) + "\n").encode()
    if _REAL_STDOUT is None:
        sys.stdout.write(data.decode())
        sys.stdout.flush()
    else:
        os.write(_REAL_STDOUT, data)


if __name__ == "__main__":
    args = parse()
    sys.stdout.flush()
    _REAL_STDOUT = os.dup(1)
    os.dup2(2, 1)
    if args.impl == "reference":
        run_reference(args)
    else:
        run_b200(args)
