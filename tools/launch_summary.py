"""Aggregate an `ncu --metrics gpu__time_duration.sum --csv` launch list by kernel name."""
import collections
import csv
import re
import sys

rows = [l for l in open(sys.argv[1]) if not l.startswith("==")]
agg = collections.OrderedDict()
for r in csv.DictReader(rows):
    if r.get("Metric Name") != "gpu__time_duration.sum":
        continue
    v = float(r["Metric Value"].replace(",", ""))
    u = r["Metric Unit"]
    v = v / 1000 if u == "ns" else v * 1000 if u == "ms" else v
    name = re.sub(r"\(.*", "", r["Kernel Name"])[:80]
    a = agg.setdefault(name, [0, 0.0])
    a[0] += 1
    a[1] += v
tot = sum(a[1] for a in agg.values())
for k, a in sorted(agg.items(), key=lambda x: -x[1][1]):
    print(f"{a[1]:10.1f} us {a[0]:4d} {100 * a[1] / tot:5.1f}% avg {a[1] / a[0]:8.1f}  {k}")
print(f"total {tot:.1f} us")
