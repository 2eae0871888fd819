#!/usr/bin/env python
"""Key metrics per kernel from `ncu --page raw --csv` exports (made on the GPU box; the .ncu-rep files are too large to
bring back):   python tools/ncu_raw_summary.py profiles/r02_*.raw.csv"""
import csv
import sys

WANT = ["gpu__time_duration.sum", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__pipe_tensor_subpipe_hmma_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "dram__throughput.avg.pct_of_peak_sustained_elapsed", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "lts__t_sector_hit_rate.pct", "lts__t_bytes.sum", "l1tex__t_sector_hit_rate.pct", "launch__registers_per_thread",
        "launch__grid_size", "launch__block_size", "launch__shared_mem_per_block_dynamic", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "smsp__cycles_active.avg", "sm__cycles_elapsed.max", "smsp__inst_executed.sum"]
for path in sys.argv[1:]:
    rows = list(csv.reader(open(path)))
    try:
        hi = next(i for i, r in enumerate(rows) if "Kernel Name" in r)
    except StopIteration:
        print(f"== {path}: no kernels"); continue
    h, units = rows[hi], rows[hi + 1]
    print(f"==== {path}")
    for r in rows[hi + 2:]:
        if len(r) < len(h):
            continue
        print("  --", r[h.index("Kernel Name")][:100], "grid", r[h.index("Grid Size")] if "Grid Size" in h else "")
        for name in WANT:
            if name in h:
                i = h.index(name)
                print(f"     {name:80s} {r[i]:>20s} {units[i]}")
