#!/usr/bin/env python
"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list per kernel.

  python tools/launch_summary.py gpurun_out/launches.csv > profiles/rNN_launch_summary.txt
"""
import collections
import csv
import sys

with open(sys.argv[1]) as f:
    lines = [l for l in f if not l.startswith("==")]
agg = collections.defaultdict(list)
for row in csv.DictReader(lines):
    agg[row["Kernel Name"].split("(")[0][:72]].append(float(row["Metric Value"].replace(",", "")))
total = sum(sum(v) for v in agg.values())
print("%-72s %5s %11s %10s %10s %6s" % ("kernel", "n", "sum_us", "mean_us", "min_us", "share"))
for k, v in sorted(agg.items(), key=lambda kv: -sum(kv[1])):
    print("%-72s %5d %11.1f %10.2f %10.2f %5.1f%%" % (k, len(v), sum(v) / 1e3, sum(v) / len(v) / 1e3, min(v) / 1e3,
                                                     100 * sum(v) / total))
