#!/usr/bin/env python
"""Per-kernel totals and shares from an ncu launch list (`--metrics gpu__time_duration.sum --csv`).
Usage: tools/launch_shares.py gpurun_out/launches.csv > profiles/x_launch_shares.txt"""
import csv, sys, collections
rows = [l for l in open(sys.argv[1]) if l.startswith('"')]
t = collections.OrderedDict()
for r in csv.DictReader(rows):
    if r["Metric Name"] != "gpu__time_duration.sum":
        continue
    name = r["Kernel Name"].split("(")[0]
    ns = float(r["Metric Value"].replace(",", "")) * {"ns": 1, "us": 1e3, "usecond": 1e3, "ms": 1e6}.get(r["Metric Unit"], 1)
    t.setdefault(name, []).append(ns)
tot = sum(sum(v) for v in t.values())
for name, v in sorted(t.items(), key=lambda kv: -sum(kv[1])):
    print(f"{name[:64]:64s} n={len(v):4d} sum={sum(v) / 1e3:10.1f} us share={sum(v) / tot:.3f} mean={sum(v) / len(v) / 1e3:.1f} us")
