"""Hottest SASS instructions of one captured launch.
usage: ncu -i X.ncu-rep --page source --csv --print-source sass --launch-count 1 | python tools/ncu_hot.py [N]"""
import csv, sys
N = int(sys.argv[1]) if len(sys.argv) > 1 else 40
rows = list(csv.reader(sys.stdin))
hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
h = rows[hi]
ci = {n: i for i, n in enumerate(h)}
data = [r for r in rows[hi + 1:] if len(r) == len(h) and r[ci["# Samples"]].isdigit()]
tot = sum(int(r[ci["# Samples"]]) for r in data)
print("kernel:", rows[0][1][:100] if rows[0] else "")
print("total samples", tot, "instructions", len(data))
idx = {id(r): k for k, r in enumerate(data)}
for r in sorted(data, key=lambda r: -int(r[ci["# Samples"]]))[:N]:
    print(f"#{idx[id(r)]:5d} {int(r[ci['# Samples']]) / tot * 100:5.2f}%  lsb {r[ci['stall_long_sb']]:>6} wait {r[ci['stall_wait']]:>5} ssb {r[ci['stall_short_sb']]:>5} "
          f"thr {r[ci['Avg. Threads Executed']]:>5} exec {r[ci['Instructions Executed']]:>9}  {r[ci['Source']][:80]}")
