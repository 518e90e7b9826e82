"""Condense `ncu -i X.ncu-rep --page raw --csv` into the handful of counters DESIGN.md argues from.

usage: ncu -i prof.ncu-rep --page raw --csv | python tools/ncu_summary.py > profiles/<name>.txt
One column per captured launch.
"""
import csv
import re
import sys

KEEP = re.compile(
    r"^(gpu__time_duration\.sum|dram__bytes_(read|write)\.sum(\.per_second)?|gpu__dram_throughput\.avg\.pct|"
    r"lts__t_sector_hit_rate\.pct|lts__t_bytes\.sum(\.per_second)?|lts__throughput\.avg\.pct|l1tex__t_sector_hit_rate\.pct|"
    r"l1tex__throughput\.avg\.pct_of_peak_sustained_active|l1tex__t_bytes\.sum|l1tex__[a-z0-9_]+\.(avg|sum)(\.pct_of_peak_sustained_(active|elapsed))?$|"
    r"lts__[a-z0-9_]+\.(avg|sum)\.pct_of_peak_sustained_elapsed$|smsp__inst_executed_op_(shared|global|local)_(ld|st)\.sum|"
    r"launch__(grid_size|block_size|registers_per_thread|occupancy_limit_registers)|"
    r"sm__warps_active\.avg\.pct|sm__throughput\.avg\.pct|smsp__issue_active\.avg\.pct|"
    r"smsp__inst_executed\.sum|smsp__thread_inst_executed_per_inst_executed\.ratio|"
    r"sm__inst_executed_pipe_(fp64|fma|alu|lsu|xu)\.avg\.pct_of_peak_sustained_active|"
    r"sass__inst_executed_local_(loads|stores)|"
    r"smsp__average_warps_issue_stalled_[a-z_]+_per_issue_active\.ratio|"
    r"smsp__average_warp_latency_per_inst_issued\.ratio)"
)

rows = list(csv.reader(sys.stdin))
hdr = None
for i, r in enumerate(rows):
    if "Kernel Name" in r:
        hdr = i
        break
if hdr is None:
    sys.exit("no header row found")
names, units, data = rows[hdr], rows[hdr + 1], [r for r in rows[hdr + 2:] if len(r) == len(rows[hdr])]
kcol = names.index("Kernel Name")
print("kernels:", [d[kcol][:60] for d in data])
for c, n in enumerate(names):
    if KEEP.match(n):
        print(f"{n:<92} {units[c]:<14} {[d[c] for d in data]}")
