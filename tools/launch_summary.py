#!/usr/bin/env python
"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list: per-kernel count, total, share."""
import csv, sys, collections
rows = []
with open(sys.argv[1]) as f:
    lines = [l for l in f if not l.startswith("==")]
rd = csv.DictReader(lines)
tot = collections.OrderedDict()
for r in rd:
    if r.get("Metric Name") != "gpu__time_duration.sum":
        continue
    name = r["Kernel Name"].split("(")[0][:70]
    v = float(r["Metric Value"].replace(",", ""))
    unit = r["Metric Unit"]
    v_us = v / 1e3 if unit in ("ns", "nsecond") else (v if unit in ("us", "usecond") else v * 1e3)
    t = tot.setdefault(name, [0, 0.0])
    t[0] += 1; t[1] += v_us
s = sum(t[1] for t in tot.values())
print(f"# total {s:.1f} us over {sum(t[0] for t in tot.values())} launches (cold-cache, serialised: shares, not absolutes)")
for name, (n, us) in sorted(tot.items(), key=lambda kv: -kv[1][1]):
    print(f"{us:12.1f} us {100*us/s:5.1f}%  x{n:4d}  avg {us/n:9.1f} us  {name}")
