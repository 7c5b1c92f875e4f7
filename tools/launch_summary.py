#!/usr/bin/env python
"""Summarise an ncu `--metrics gpu__time_duration.sum --csv` launch list: per-kernel count, total and share.
usage: launch_summary.py launches.csv [n_steps_in_capture]"""
import csv, re, sys
from collections import defaultdict
path = sys.argv[1]
steps = float(sys.argv[2]) if len(sys.argv) > 2 else 1.0
rows = []
with open(path) as f:
    lines = [l for l in f if l.startswith('"')]
rd = csv.DictReader(lines)
agg = defaultdict(lambda: [0, 0.0])
for r in rd:
    if r["Metric Name"] != "gpu__time_duration.sum":
        continue
    name = r["Kernel Name"]
    name = re.sub(r"\(.*$", "", name)
    name = re.sub(r"^void ", "", name)
    if len(name) > 90:
        name = name[:90]
    v = float(r["Metric Value"].replace(",", ""))
    if r["Metric Unit"] == "us": v *= 1e3
    elif r["Metric Unit"] == "ms": v *= 1e6
    a = agg[name]; a[0] += 1; a[1] += v
tot = sum(a[1] for a in agg.values())
print(f"total {tot/1e6/steps:.3f} ms per step over {steps:g} steps; {sum(a[0] for a in agg.values())/steps:.0f} launches per step")
for name, (n, t) in sorted(agg.items(), key=lambda kv: -kv[1][1])[:45]:
    print(f"{t/1e6/steps:9.3f} ms {100*t/tot:5.1f}%  n={n/steps:7.1f}  avg {t/n/1e3:9.1f} us  {name}")
