#!/usr/bin/env python3
"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list per kernel."""
import collections
import csv
import sys

lines = [l for l in open(sys.argv[1]) if not l.startswith("==")]
agg = collections.OrderedDict()
for row in csv.DictReader(lines):
    v = float(row["Metric Value"].replace(",", ""))
    u = row["Metric Unit"]
    ns = v if u in ("ns", "nsecond") else v * 1e3 if u in ("us", "usecond") else v * 1e6 if u in ("ms", "msecond") else v
    agg.setdefault(row["Kernel Name"].split("(")[0][:64], []).append(ns)
tot = sum(sum(v) for v in agg.values())
print(f"{'kernel':66s} {'n':>4s} {'avg us':>9s} {'share':>7s}")
for k, v in agg.items():
    print(f"{k:66s} {len(v):4d} {sum(v) / len(v) / 1e3:9.2f} {100 * sum(v) / tot:6.1f}%")
