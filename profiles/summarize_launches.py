#!/usr/bin/env python
"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list: per-kernel count, average us, share."""
import csv, sys, collections
path = sys.argv[1]
rows = [l for l in open(path) if l.startswith('"')]
rd = csv.DictReader(rows)
agg = collections.OrderedDict()
for r in rd:
    if r.get("Metric Name") != "gpu__time_duration.sum":
        continue
    name = r["Kernel Name"].split("(")[0]
    v = float(r["Metric Value"].replace(",", ""))
    unit = r["Metric Unit"]
    us = v / 1e3 if unit in ("ns", "nsecond") else (v if unit in ("us", "usecond") else v * 1e3)
    a = agg.setdefault(name, [0, 0.0]); a[0] += 1; a[1] += us
tot = sum(a[1] for a in agg.values())
print(f"# {path}: {sum(a[0] for a in agg.values())} launches, total {tot/1e3:.3f} ms (gpu__time_duration.sum, cold-cache serialised: compare SHARES)")
print(f"{'kernel':70s} {'n':>6s} {'avg_us':>10s} {'share%':>7s}")
for k, a in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print(f"{k[:70]:70s} {a[0]:6d} {a[1]/a[0]:10.2f} {100*a[1]/tot:7.2f}")
