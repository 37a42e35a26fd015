#!/usr/bin/env python
"""Per-kernel launch count and mean / min duration from an `ncu --metrics gpu__time_duration.sum --csv` log."""
import collections
import csv
import sys

lines = [l for l in open(sys.argv[1]) if not l.startswith("==")]
agg = collections.defaultdict(list)
for row in csv.DictReader(lines):
    if row.get("Metric Name") == "gpu__time_duration.sum":
        v = float(row["Metric Value"].replace(",", ""))
        agg[row["Kernel Name"][:70]].append(v / 1000 if row["Metric Unit"] in ("ns", "nsecond") else v)
for k, v in agg.items():
    print(f"{k:72s} n={len(v):3d} mean={sum(v) / len(v):9.1f} us  min={min(v):8.1f}")
