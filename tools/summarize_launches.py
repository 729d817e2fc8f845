#!/usr/bin/env python
"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list: per-kernel count / total / avg / share."""
import csv, sys, collections
path = sys.argv[1]
skip = int(sys.argv[2]) if len(sys.argv) > 2 else 0      # launches to skip (setup / warm-up)
rows = []
with open(path) as f:
    lines = [l for l in f if not l.startswith("==")]
rd = csv.DictReader(lines)
for r in rd:
    if r.get("Metric Name") != "gpu__time_duration.sum":
        continue
    v = float(r["Metric Value"].replace(",", ""))
    unit = r["Metric Unit"]
    us = v / 1000.0 if unit in ("ns", "nsecond") else (v if unit in ("us", "usecond") else v * 1000.0)
    rows.append((r["Kernel Name"], us))
rows = rows[skip:]
agg = collections.OrderedDict()
for k, us in rows:
    k = k[:60]
    a = agg.setdefault(k, [0, 0.0]); a[0] += 1; a[1] += us
tot = sum(a[1] for a in agg.values())
print(f"| kernel | launches | total us | avg us | share |\n|---|---:|---:|---:|---:|")
for k, a in sorted(agg.items(), key=lambda kv: -kv[1][1])[:30]:
    print(f"| `{k}` | {a[0]} | {a[1]:.1f} | {a[1]/a[0]:.2f} | {100*a[1]/tot:.1f}% |")
print(f"total {tot:.1f} us over {len(rows)} launches")
