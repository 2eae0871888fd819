"""Key metrics of every kernel in an .ncu-rep (read on the CPU box):  python tools/ncu_summary.py REP [more metric substrings]"""
import csv, subprocess, sys
rep = sys.argv[1]
extra = sys.argv[2:]
out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
h, units = rows[0], rows[1]
want = ["gpu__time_duration.sum", "sm__cycles_elapsed.max", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_uniform", "dram__bytes_read.sum", "dram__bytes_write.sum", "lts__t_sector_hit_rate.pct",
        "lts__t_sectors_srcunit_tex.sum", "launch__registers_per_thread", "launch__grid_size", "smsp__inst_executed.sum",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "smsp__cycles_active.avg", "sm__cycles_active.avg",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum",
        "smsp__average_warp_latency_issue_stalled", "smsp__average_warps_issue_stalled"] + extra
for r in rows[2:]:
    print("==", r[h.index("Kernel Name")][:90])
    for i, name in enumerate(h):
        if any(name == w or (w in extra and w in name) for w in want) or any(name.startswith(w) for w in want[15:]):
            print(f"  {name:90s} {r[i]:>18s} {units[i]}")
