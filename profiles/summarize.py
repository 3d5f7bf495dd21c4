#!/usr/bin/env python
"""Turn an .ncu-rep (ncu --set full --import-source on) into the short text summary kept under profiles/.

    python profiles/summarize.py gpurun_out/X.ncu-rep > profiles/X_summary.txt

Reads the report with `ncu -i ... --page raw --csv` and `--page source --csv` (no GPU needed)."""
import collections
import csv
import io
import subprocess
import sys

KEYS = [
    "gpu__time_duration.sum", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
    "launch__shared_mem_per_block_dynamic", "launch__shared_mem_per_block_static", "launch__occupancy_limit_registers",
    "launch__occupancy_limit_shared_mem", "launch__occupancy_limit_warps", "sm__warps_active.avg.pct_of_peak_sustained_active",
    "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
    "lts__throughput.avg.pct_of_peak_sustained_elapsed", "l1tex__throughput.avg.pct_of_peak_sustained_active",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed", "smsp__issue_active.avg.pct_of_peak_sustained_active",
    "smsp__inst_executed.sum", "smsp__thread_inst_executed_per_inst_executed.ratio",
    "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_tensor_subpipe_hmma.avg.pct_of_peak_sustained_active",
    "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum",
    "sm__cycles_elapsed.max", "sm__cycles_elapsed.max.per_second",
]


def ncu(path, page, extra=()):
    return subprocess.run(["ncu", "-i", path, "--page", page, "--csv", *extra], capture_output=True, text=True).stdout


def main():
    path = sys.argv[1]
    top = int(sys.argv[2]) if len(sys.argv) > 2 else 30
    rows = list(csv.reader(io.StringIO(ncu(path, "raw"))))
    hdr, units, data = rows[0], rows[1], rows[2:]
    name_col = hdr.index("Kernel Name")
    print("# %s" % path)
    for r in data:
        print("\n## launch id %s: %s" % (r[0], r[name_col][:120]))
        for i, h in enumerate(hdr):
            if h in KEYS or (h.startswith("smsp__average_warps_issue_stalled") and h.endswith("per_issue_active.ratio")
                             and float(r[i] or 0) >= 0.2):
                print("%-86s %-14s %s" % (h, units[i], r[i]))
    src = list(csv.reader(io.StringIO(ncu(path, "source", ("--print-source", "cuda,sass")))))
    cur, h2, agg = None, None, collections.OrderedDict()
    for r in src:
        if not r:
            continue
        if r[0] == "File Path":
            cur, h2 = r[1].split("/")[-1], None
            continue
        if r[0] == "Line No":
            h2 = r
            continue
        if h2 is None:
            continue
        d = dict(zip(h2, r))
        try:
            ln, ie = int(d["Line No"]), int(d.get("Instructions Executed") or 0)
            te, s = int(d.get("Thread Instructions Executed") or 0), int(d.get("# Samples") or 0)
        except (ValueError, KeyError):
            continue
        if ie > 0:
            k = (cur, ln, r[1].strip()[:96])
            a = agg.get(k, (0, 0, 0))
            agg[k] = (a[0] + ie, a[1] + te, a[2] + s)
    tot = sum(v[0] for v in agg.values()) or 1
    byfile = collections.Counter()
    for k, v in agg.items():
        byfile[k[0]] += v[0]
    print("\n## warp-instructions by source file (all profiled launches): " +
          ", ".join("%s %.1f%%" % (k, 100.0 * v / tot) for k, v in byfile.most_common(8)))
    print("## top source lines: share of warp-instructions | avg active threads | stall samples | file:line | source")
    for k, v in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
        print("%5.2f%% | %4.1f | %6d | %s:%d | %s" % (100.0 * v[0] / tot, v[1] / max(v[0], 1), v[2], k[0], k[1], k[2]))


if __name__ == "__main__":
    main()
