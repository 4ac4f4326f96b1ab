#!/usr/bin/env python
"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list: time share per kernel."""
import csv
import re
import sys
from collections import defaultdict

rows = []
with open(sys.argv[1]) as f:
    lines = [l for l in f if l.startswith('"')]
for r in csv.DictReader(lines):
    if r.get("Metric Name") == "gpu__time_duration.sum":
        rows.append((r["Kernel Name"], float(r["Metric Value"]), r["Grid Size"]))
tot = sum(t for _, t, _ in rows)
agg = defaultdict(lambda: [0, 0.0])
for k, t, g in rows:
    k = re.sub(r"^void ", "", k)
    k = k.replace("(int)", "")
    k = re.sub(r"\(.*$", "", k)
    agg[k][0] += 1
    agg[k][1] += t
print("launches %d   total device time %.3f ms" % (len(rows), tot / 1e6))
print("%-70s %7s %10s %7s %9s" % ("kernel", "count", "ms", "share", "avg us"))
for k, (c, t) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print("%-70s %7d %10.3f %6.1f%% %9.1f" % (k[:70], c, t / 1e6, 100 * t / tot, t / c / 1e3))
