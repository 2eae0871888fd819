"""Per-kernel DRAM traffic and bandwidth out of an ncu csv with gpu__time_duration.sum, dram__bytes_read.sum,
dram__bytes_write.sum, lts__t_sector_hit_rate.pct:   python tools/dram_table.py FILE.csv [merge]
`merge` adds up launches of the same kernel name."""
import csv, sys
from collections import OrderedDict
rows = list(csv.reader(open(sys.argv[1])))
merge = len(sys.argv) > 2
hi = next(i for i, r in enumerate(rows) if "Kernel Name" in r)
h = rows[hi]
ki, mi, ui, vi, ii = h.index("Kernel Name"), h.index("Metric Name"), h.index("Metric Unit"), h.index("Metric Value"), h.index("ID")
SCALE = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "ns": 1e-9, "us": 1e-6, "ms": 1e-3, "s": 1.0, "%": 1.0, "usecond": 1e-6,
         "nsecond": 1e-9, "msecond": 1e-3, "second": 1.0}
launches = OrderedDict()
for r in rows[hi + 1:]:
    if len(r) <= vi:
        continue
    key = r[ii]
    d = launches.setdefault(key, {"name": r[ki].split("(")[0][:48]})
    try:
        d[r[mi]] = float(r[vi].replace(",", "")) * SCALE.get(r[ui], 1.0)
    except ValueError:
        pass
out = OrderedDict()
for d in launches.values():
    t = d.get("gpu__time_duration.sum", 0.0)
    rd, wr = d.get("dram__bytes_read.sum", 0.0), d.get("dram__bytes_write.sum", 0.0)
    k = d["name"] if merge else (d["name"], len(out))
    o = out.setdefault(k, {"name": d["name"], "n": 0, "t": 0.0, "rd": 0.0, "wr": 0.0, "hit": 0.0})
    o["n"] += 1; o["t"] += t; o["rd"] += rd; o["wr"] += wr; o["hit"] += d.get("lts__t_sector_hit_rate.pct", 0.0)
print(f"{'kernel':48s} {'launches':>8s} {'time us':>10s} {'DRAM rd MB':>11s} {'DRAM wr MB':>11s} {'DRAM GB/s':>10s} {'L2 hit %':>9s}")
for o in out.values():
    gbps = (o["rd"] + o["wr"]) / o["t"] / 1e9 if o["t"] > 0 else 0.0
    print(f"{o['name']:48s} {o['n']:8d} {o['t'] * 1e6:10.1f} {o['rd'] / 1e6:11.2f} {o['wr'] / 1e6:11.2f} {gbps:10.1f} {o['hit'] / o['n']:9.1f}")
