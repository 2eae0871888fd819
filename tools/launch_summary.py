"""Per-kernel times of ONE hot-path invocation out of an ncu launch list (gpu__time_duration.sum csv):
    python tools/launch_summary.py LAUNCHES.csv [which]
`which` = index of the invocation to print (each starts at split_rows_kernel), default the last full one."""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
hi = next(i for i, r in enumerate(rows) if "Kernel Name" in r)
h = rows[hi]
ki, vi = h.index("Kernel Name"), h.index("Metric Value")
names, vals = [], []
for r in rows[hi + 1:]:
    if len(r) > vi:
        try:
            vals.append(float(r[vi].replace(",", ""))); names.append(r[ki])
        except ValueError:
            pass
starts = [i for i, n in enumerate(names) if "split_rows" in n]
which = int(sys.argv[2]) if len(sys.argv) > 2 else len(starts) - 2
s, e = starts[which], starts[which + 1]
tot = 0.0
for i in range(s, e):
    print(f"{names[i][:64]:64s} {vals[i] / 1e3:9.1f} us"); tot += vals[i]
print(f"{'sum of kernel durations':64s} {tot / 1e6:9.3f} ms  ({e - s} launches)")
