"""Aggregate an `ncu --metrics gpu__time_duration.sum --csv` launch list: per kernel count / mean / total (us)."""
import csv
import sys
from collections import OrderedDict

rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 10]
hdr = rows[0]
ki, vi, ui = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
agg = OrderedDict()
for r in rows[1:]:
    v = float(r[vi].replace(",", ""))
    v = {"ns": v / 1e3, "us": v, "ms": v * 1e3, "nsecond": v / 1e3, "usecond": v, "msecond": v * 1e3}.get(r[ui], v)
    agg.setdefault(r[ki][:70], []).append(v)
for k, v in agg.items():
    print("%-72s n=%4d mean=%9.1f us total=%10.1f us" % (k, len(v), sum(v) / len(v), sum(v)))
