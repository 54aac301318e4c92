#!/usr/bin/env python
"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list per kernel (share of GPU time)."""
import collections
import csv
import sys

rows = [r for r in csv.reader(l for l in open(sys.argv[1]) if not l.startswith("=="))]
hdr = rows[0]
ki, vi = hdr.index("Kernel Name"), hdr.index("Metric Value")
agg = collections.defaultdict(list)
for r in rows[1:]:
    if len(r) > vi:
        try:
            agg[r[ki].split("(")[0][-70:]].append(float(r[vi].replace(",", "")))
        except ValueError:
            pass
tot = sum(sum(v) for v in agg.values())
print("%-72s %5s %12s %10s %7s" % ("kernel", "n", "sum_us", "avg_us", "share"))
for k, v in sorted(agg.items(), key=lambda kv: -sum(kv[1])):
    print("%-72s %5d %12.1f %10.2f %7.3f" % (k, len(v), sum(v) / 1e3, sum(v) / len(v) / 1e3, sum(v) / tot))
