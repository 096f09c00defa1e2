#!/usr/bin/env python
"""Aggregate an `ncu --metrics gpu__time_duration.sum --csv` launch list by kernel name.

usage: python scripts/summarize_launches.py gpurun_out/launches.csv [--last N] [--out profiles/x.md]
"""
import csv
import io
import re
import sys
from collections import defaultdict

path = sys.argv[1]
last = int(sys.argv[sys.argv.index("--last") + 1]) if "--last" in sys.argv else None
out = sys.argv[sys.argv.index("--out") + 1] if "--out" in sys.argv else None
lines = open(path, errors="replace").read().splitlines()
start = next(i for i, l in enumerate(lines) if l.startswith('"ID"'))
rows = list(csv.DictReader(io.StringIO("\n".join(lines[start:]))))
rows = [r for r in rows if r.get("Metric Name") == "gpu__time_duration.sum"]
if last:
    rows = rows[-last:]
agg = defaultdict(lambda: [0, 0.0])
tot = 0.0
for r in rows:
    name = re.sub(r"\(.*", "", r["Kernel Name"])
    name = re.sub(r"^void ", "", name)
    v = float(r["Metric Value"].replace(",", ""))
    unit = r.get("Metric Unit", "ns")
    us = v / 1000.0 if unit in ("ns", "nsecond") else (v if unit in ("us", "usecond") else v * 1000.0)
    agg[name][0] += 1
    agg[name][1] += us
    tot += us
text = [f"launches: {len(rows)}  total device time: {tot / 1000:.3f} ms (serialised, cold caches: compare shares)",
        "", "| kernel | launches | total us | avg us | share |", "|---|---:|---:|---:|---:|"]
for name, (n, us) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    text.append(f"| `{name[:90]}` | {n} | {us:.1f} | {us / n:.2f} | {100 * us / tot:.1f}% |")
s = "\n".join(text)
print(s)
if out:
    open(out, "w").write(s + "\n")
