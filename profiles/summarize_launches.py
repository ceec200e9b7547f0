"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list: per-kernel launches,
total device time and share.  Usage: python profiles/summarize_launches.py launches.csv"""
import csv
import re
import sys
from collections import defaultdict

rows = []
with open(sys.argv[1]) as f:
    lines = [l for l in f if not l.startswith("==")]
rd = csv.DictReader(lines)
tot = defaultdict(lambda: [0, 0.0])
for r in rd:
    if r.get("Metric Name") != "gpu__time_duration.sum":
        continue
    name = re.sub(r"\(.*", "", r["Kernel Name"])
    v = float(r["Metric Value"].replace(",", ""))
    unit = r.get("Metric Unit", "ns")
    v = v * {"ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6}.get(unit, 1e-3)
    tot[name][0] += 1
    tot[name][1] += v
all_us = sum(v[1] for v in tot.values())
print("%-60s %8s %12s %8s %10s" % ("kernel", "launches", "total_us", "share", "avg_us"))
for k, (n, us) in sorted(tot.items(), key=lambda kv: -kv[1][1]):
    print("%-60s %8d %12.1f %7.1f%% %10.2f" % (k[:60], n, us, 100 * us / all_us, us / n))
print("%-60s %8d %12.1f" % ("TOTAL", sum(v[0] for v in tot.values()), all_us))
