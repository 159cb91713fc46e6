"""Summarise an ncu launch list (`ncu --metrics gpu__time_duration.sum --csv --log-file X.csv ...`): launches grouped
by kernel and grid, with count, average and total duration.  usage: python tools/ncu_launches.py X.csv ["command"]"""
import collections
import csv
import sys

rows = [r for r in csv.reader(open(sys.argv[1], errors="replace")) if len(r) > 10]
hdr = rows[0]
col = {h: i for i, h in enumerate(hdr)}
agg = collections.OrderedDict()
for r in rows[1:]:
    if r[col["Metric Name"]] != "gpu__time_duration.sum":
        continue
    v = float(r[col["Metric Value"]].replace(",", ""))
    unit = r[col["Metric Unit"]]
    us = v / 1e3 if unit in ("ns", "nsecond") else (v * 1e3 if unit in ("ms", "msecond") else v)
    key = (r[col["Kernel Name"]][:110], r[col["Grid Size"]])
    a = agg.setdefault(key, [0, 0.0])
    a[0] += 1
    a[1] += us
tot = sum(a[1] for a in agg.values())
if len(sys.argv) > 2:
    print(sys.argv[2])
print("launches grouped by kernel and grid (cold-cache, serialised durations; setup kernels of the ring construction included)")
for (name, grid), (n, us) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print(f"{n:5d} x avg {us / n:9.2f} us  total {us / 1e3:8.3f} ms ({100 * us / tot:5.1f}%)  {name} grid {grid}")
