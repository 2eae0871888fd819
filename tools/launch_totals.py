"""Totals per kernel name over a whole ncu launch list (gpu__time_duration.sum csv):
    python tools/launch_totals.py LAUNCHES.csv"""
import csv, sys
from collections import OrderedDict
rows = list(csv.reader(open(sys.argv[1])))
hi = next(i for i, r in enumerate(rows) if "Kernel Name" in r)
h = rows[hi]
ki, vi = h.index("Kernel Name"), h.index("Metric Value")
tot = OrderedDict()
for r in rows[hi + 1:]:
    if len(r) > vi:
        try:
            v = float(r[vi].replace(",", ""))
        except ValueError:
            continue
        name = r[ki].split("(")[0][:60]
        t = tot.setdefault(name, [0, 0.0])
        t[0] += 1; t[1] += v
all_ns = sum(t[1] for t in tot.values())
for name, (cnt, ns) in sorted(tot.items(), key=lambda kv: -kv[1][1]):
    print(f"{name:60s} {cnt:6d} launches {ns / 1e6:9.3f} ms  {ns / cnt / 1e3:8.1f} us each  {100 * ns / all_ns:5.1f} %")
print(f"{'total':60s} {sum(t[0] for t in tot.values()):6d} launches {all_ns / 1e6:9.3f} ms")
