#!/usr/bin/env python
"""Per-kernel ncu figures of THIS library build -> profiles/traffic.json, keyed by the hash of the sources it was built
from (build.source_sha(): nvcc's output is not byte-reproducible, so the .so's own SHA-1 -- recorded too -- changes with
every rebuild of the same code).

    # on the GPU box (tools/profile_round.sh does this), after the plain bench run has exited 0:
    ncu --metrics $(python tools/update_traffic.py --metrics) --clock-control none -k regex:greedy_kernel -s 3 -c 4 \
        --csv --log-file gpurun_out/traffic_headline.csv python bench.py --steps 4 --warmup 3 --no-cpu-baseline --no-parity --no-configs --no-full-path
    python tools/update_traffic.py gpurun_out/traffic_headline.csv headline
    (and the same with -k regex:nw_band_kernel, without --no-full-path, label full_path)

bench.py prints roofline.traffic / roofline.issue only when the entry's source hash equals that of the tree the loaded
library was built from, so the figures can never silently belong to other code.  One entry per (kernel template, grid size): the median launch."""
import csv
import hashlib
import json
import os
import re
import statistics
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
METRICS = ("dram__bytes_read.sum,dram__bytes_write.sum,smsp__inst_executed.sum,"
           "smsp__issue_active.avg.pct_of_peak_sustained_active,sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active,"
           "smsp__thread_inst_executed_per_inst_executed.ratio,gpu__time_duration.sum")


def lib_sha(path):
    h = hashlib.sha1()
    with open(path, "rb") as f:
        for blk in iter(lambda: f.read(1 << 20), b""):
            h.update(blk)
    return h.hexdigest()[:16]


def to_bytes(v, unit):
    return float(v.replace(",", "")) * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(unit, 1)


def main():
    if len(sys.argv) > 1 and sys.argv[1] == "--metrics":
        print(METRICS)
        return
    rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 10]
    hdr = rows[0]
    col = {h: i for i, h in enumerate(hdr)}
    per = {}
    for r in rows[1:]:
        name = re.sub(r"\(.*", "", r[col["Kernel Name"]]).replace("void ", "").replace("gasma::", "")
        key = "%s|grid=%s" % (name, r[col["Grid Size"]].replace(",", "").split()[0] if "Grid Size" in col else "?")
        d = per.setdefault((key, r[col["ID"]]), {})
        v, unit = r[col["Metric Value"]], r[col["Metric Unit"]]
        m = r[col["Metric Name"]]
        d[m] = to_bytes(v, unit) if m.startswith("dram__bytes") else float(v.replace(",", ""))
    kernels = {}
    for (key, _id), d in per.items():
        kernels.setdefault(key, []).append(d)
    lib = os.path.join(ROOT, "approximate-string-matching_b200", "libgasma_b200.so")
    label = sys.argv[2] if len(sys.argv) > 2 else "headline"
    sha = lib_sha(lib)
    import importlib.util
    spec = importlib.util.spec_from_file_location("gasma_b200_build", os.path.join(ROOT, "approximate-string-matching_b200", "build.py"))
    gasma_build = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(gasma_build)
    src = gasma_build.source_sha()
    path = os.path.join(ROOT, "profiles", "traffic.json")
    try:
        with open(path) as f:
            entry = json.load(f)
    except Exception:
        entry = {}
    if entry.get("source_sha1") != src:       # figures of other sources are dropped, never mixed
        entry = {"library_sha1": sha, "source_sha1": src, "legs": {}}
    key, ls = max(kernels.items(), key=lambda kv: len(kv[1]))     # the leg's dominant kernel = the one launched most
    med = lambda m: statistics.median(x[m] for x in ls if m in x)
    entry["legs"][label] = {
        "kernel": key, "launches": len(ls), "source": os.path.basename(sys.argv[1]),
        "traffic_bytes": med("dram__bytes_read.sum") + med("dram__bytes_write.sum"),
        "dram_read_bytes": med("dram__bytes_read.sum"), "dram_write_bytes": med("dram__bytes_write.sum"),
        "warp_instructions": med("smsp__inst_executed.sum"),
        "issue_active_pct": med("smsp__issue_active.avg.pct_of_peak_sustained_active"),
        "alu_pipe_pct": med("sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active"),
        "avg_active_threads": med("smsp__thread_inst_executed_per_inst_executed.ratio"),
        "ncu_time_us": med("gpu__time_duration.sum"),
    }
    with open(path, "w") as f:
        json.dump(entry, f, indent=1, sort_keys=True)
    print(path, sha, label, key)


if __name__ == "__main__":
    main()
