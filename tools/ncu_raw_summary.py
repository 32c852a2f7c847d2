"""Distil `ncu -i X.ncu-rep --page raw --csv` (one row per captured launch) into the text summaries kept under profiles/: the
launch, occupancy, issue, pipe and DRAM counters plus the top stall reasons per issue of every launch.
usage: python tools/ncu_raw_summary.py RAW.csv "header line" > profiles/...txt"""
import csv
import re
import sys

rows = list(csv.reader(open(sys.argv[1])))
print(sys.argv[2])
hdr, units = rows[0], rows[1]
keep = ["gpu__time_duration.sum", "launch__registers_per_thread", "launch__grid_size", "launch__block_size", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "smsp__issue_active.avg.per_cycle_active", "smsp__inst_executed.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "sm__pipe_fmaheavy_cycles_active.avg.pct_of_peak_sustained_elapsed",
        "sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active", "lts__t_sector_hit_rate.pct"]
for r in rows[2:]:
    d = dict(zip(hdr, r))
    u = dict(zip(hdr, units))
    print("----")
    print("Kernel Name ", d.get("Kernel Name", "?"))
    for k in keep:
        if k in d:
            print(k, u[k], d[k])
    stalls = []
    for k, v in d.items():
        m = re.match(r"smsp__average_warps_issue_stalled_(\w+)_per_issue_active\.ratio", k)
        if m and v:
            try:
                stalls.append((float(v.replace(",", "")), m.group(1)))
            except ValueError:
                pass
    for v, name in sorted(stalls, reverse=True)[:6]:
        print("   stall %-28s %6.2f" % (name, v))
