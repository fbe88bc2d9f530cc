#!/usr/bin/env python
"""Aggregates an `ncu --metrics gpu__time_duration.sum --csv` launch list per kernel."""
import collections
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
hdr, agg = None, collections.defaultdict(lambda: [0, 0.0, 0.0])
for r in rows:
    if "Kernel Name" in r:
        hdr = r
        continue
    if hdr is None or len(r) != len(hdr):
        continue
    d = dict(zip(hdr, r))
    if d.get("Metric Name") != "gpu__time_duration.sum":
        continue
    v, u = float(d["Metric Value"].replace(",", "")), d["Metric Unit"]
    ms = v / 1e6 if u.startswith("n") else v / 1e3 if u.startswith("u") else v * 1e3 if u in ("s", "second") else v
    k = d["Kernel Name"].split("(")[0]
    a = agg[k]
    a[0] += 1
    a[1] += ms
    a[2] = max(a[2], ms)
tot = sum(a[1] for a in agg.values())
print(f"{'kernel':28s} {'launches':>8s} {'total ms':>10s} {'share':>7s} {'avg ms':>9s} {'max ms':>9s}")
for k, a in sorted(agg.items(), key=lambda x: -x[1][1]):
    print(f"{k:28s} {a[0]:8d} {a[1]:10.2f} {a[1] / tot * 100:6.1f}% {a[1] / a[0]:9.3f} {a[2]:9.3f}")
print(f"{'total':28s} {sum(a[0] for a in agg.values()):8d} {tot:10.2f}")
