#!/usr/bin/env python
"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list: time share per kernel/grid."""
import collections
import csv
import re
import sys

path = sys.argv[1]
with open(path) as f:
    lines = [l for l in f if not l.startswith("==")]
agg, tot = collections.OrderedDict(), 0.0
for row in csv.DictReader(lines):
    name = re.sub(r"^.*::", "", re.sub(r"\(.*", "", row["Kernel Name"]))
    v = float(row["Metric Value"]) / 1000.0
    a = agg.setdefault((name, row.get("Grid Size", "")), [0, 0.0])
    a[0] += 1
    a[1] += v
    tot += v
print(f"total {tot:.1f} us over {sum(a[0] for a in agg.values())} launches (cold-cache, serialised: compare shares)")
for (n, g), (c, t) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print(f"{t:9.1f} us {100 * t / tot:5.1f}%  n={c:3d} avg={t / c:7.2f} us  {n[:56]:56s} grid={g}")
