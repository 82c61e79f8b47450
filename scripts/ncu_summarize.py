#!/usr/bin/env python
"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` log: per-kernel count, mean and total microseconds."""
import collections
import csv
import sys

rows = []
with open(sys.argv[1]) as f:
    lines = [l for l in f if l.startswith('"')]
rd = csv.DictReader(lines)
tot = collections.OrderedDict()
for r in rd:
    if r.get("Metric Name") != "gpu__time_duration.sum":
        continue
    name = r["Kernel Name"].split("(")[0]
    v = float(r["Metric Value"].replace(",", ""))
    unit = r.get("Metric Unit", "ns")
    us = v / 1e3 if unit in ("ns", "nsecond") else v if unit in ("us", "usecond") else v * 1e3
    d = tot.setdefault(name, [0, 0.0])
    d[0] += 1; d[1] += us
skip = int(sys.argv[2]) if len(sys.argv) > 2 else 0
print("%-60s %6s %10s %10s" % ("kernel", "count", "mean_us", "total_us"))
for k, (n, t) in sorted(tot.items(), key=lambda kv: -kv[1][1]):
    print("%-60s %6d %10.2f %10.1f" % (k[:60], n, t / n, t))
