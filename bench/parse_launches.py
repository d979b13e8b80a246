#!/usr/bin/env python
"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list: per-kernel mean/total and
one step's kernel sequence. Usage: python bench/parse_launches.py gpurun_out/launches.csv"""
import collections
import csv
import re
import sys

path = sys.argv[1]
with open(path) as f:
    lines = [l for l in f if not l.startswith("==")]
agg = collections.defaultdict(list)
order = []
for row in csv.DictReader(lines):
    try:
        v = float(row["Metric Value"].replace(",", ""))
    except (ValueError, KeyError):
        continue
    unit = row["Metric Unit"]
    v = v / 1000 if unit == "ns" else v * 1000 if unit == "ms" else v
    short = re.sub(r"\(.*", "", row["Kernel Name"])
    grid = row.get("Grid Size", "")
    agg[(short, grid)].append(v)
    order.append((short, grid, v))
for (k, g), v in sorted(agg.items(), key=lambda kv: -sum(kv[1])):
    print(f"{k[:62]:62s} grid={g:16s} n={len(v):3d} mean={sum(v)/len(v):8.2f}us")
start = next((i for i, (k, g, v) in enumerate(order) if "fill_u32" in k and i > 40), None)
if start is not None:
    print("--- one step")
    tot = 0.0
    for k, g, v in order[start:]:
        if "fill_u32" in k and tot > 0:
            break
        print(f"{k[:56]:56s} {g:16s} {v:7.2f}")
        if "at::" not in k:
            tot += v
    print(f"sum of our kernels: {tot:.1f} us")
