#!/usr/bin/env python
"""Per-kernel summary of an ncu launch list (`ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --csv`):
launches, median duration, share of the summed kernel time, median DRAM bytes.   python tools/launch_summary.py launches.csv"""
import collections
import csv
import statistics
import sys

rows = list(csv.reader(open(sys.argv[1])))
hi = [i for i, r in enumerate(rows) if "Kernel Name" in r][0]
h = rows[hi]
ki, mi, vi, ii = h.index("Kernel Name"), h.index("Metric Name"), h.index("Metric Value"), h.index("ID")
per = collections.defaultdict(dict)
for r in rows[hi + 1:]:
    if len(r) > vi:
        per[(r[ii], r[ki].split("(")[0][:64])][r[mi]] = float(r[vi].replace(",", ""))
agg = collections.defaultdict(list)
for (_, k), m in per.items():
    agg[k].append(m)
tot = sum(m.get("gpu__time_duration.sum", 0) for ms in agg.values() for m in ms)
print(f"{len(per)} launches, {tot / 1e6:.1f} ms of kernel time (per-launch times under ncu are serialised and cold-cache: read the shares)")
for k, ms in sorted(agg.items(), key=lambda x: -sum(m.get("gpu__time_duration.sum", 0) for m in x[1])):
    t = [m.get("gpu__time_duration.sum", 0) for m in ms]
    rd = statistics.median([m.get("dram__bytes_read.sum", 0) for m in ms]) / 1e9
    wr = statistics.median([m.get("dram__bytes_write.sum", 0) for m in ms]) / 1e9
    print(f"{k:66s} n={len(ms):3d}  median {statistics.median(t) / 1e6:8.3f} ms  share {sum(t) / tot * 100:5.1f} %  dram rd {rd:6.2f} GB  wr {wr:5.2f} GB")
