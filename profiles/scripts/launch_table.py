"""Aggregates an ncu launch list (--metrics gpu__time_duration.sum --csv) per (kernel, grid): launches, mean and total of
the serialised durations.    python profiles/scripts/launch_table.py launches.csv"""
import collections
import csv
import re
import sys

rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 10 and r[0].isdigit()]
agg = collections.OrderedDict()
for r in rows:
    name = re.sub(r"\(.*", "", r[4]).replace("ptts::", "").replace("(anonymous namespace)::", "").replace("void ", "")
    val, unit = float(r[-1].replace(",", "")), r[-2]
    us = val / 1000.0 if unit.startswith("n") else (val if unit.startswith("u") else val * 1000.0)
    a = agg.setdefault((name[:64], r[8]), [0, 0.0])
    a[0] += 1
    a[1] += us
tot = sum(a[1] for a in agg.values())
print(f"launches {len(rows)}, sum of serialised durations {tot / 1000:.2f} ms")
for (name, grid), (n, us) in sorted(agg.items(), key=lambda kv: -kv[1][1])[:30]:
    print(f"  {name:64s} grid {grid:14s} n={n:4d} {us / n:8.1f} us each {us / 1000:7.2f} ms")
