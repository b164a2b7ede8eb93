"""Sums an ncu `--metrics gpu__time_duration.sum --csv` launch list per kernel name.
    python profiles/scripts/ncu_launch_summary.py launches.csv [skip_first_n_launches]"""
import csv
import re
import sys
from collections import OrderedDict

rows = []
with open(sys.argv[1], newline="") as f:
    lines = [l for l in f if l.startswith('"')]
for r in csv.DictReader(lines):
    if r.get("Metric Name") == "gpu__time_duration.sum":
        val = float(r["Metric Value"].replace(",", ""))
        unit = r.get("Metric Unit", "ns")
        us = val / 1e3 if unit in ("ns", "nsecond") else val if unit in ("us", "usecond") else val * 1e3
        rows.append((re.sub(r"\(.*", "", r["Kernel Name"]), us))
skip = int(sys.argv[2]) if len(sys.argv) > 2 else 0
rows = rows[skip:]
agg = OrderedDict()
for k, us in rows:
    a = agg.setdefault(k, [0, 0.0])
    a[0] += 1
    a[1] += us
tot = sum(a[1] for a in agg.values())
print(f"{len(rows)} launches, {tot:.1f} us in total (times under ncu are serialised and cold-cache: read the shares)")
for k, (n, us) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print(f"{us:12.1f} us  {100 * us / tot:5.1f} %  x{n:<5d} {k}")
