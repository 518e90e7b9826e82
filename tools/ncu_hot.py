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

# ---- where the executed warp instructions are: consecutive SASS instructions with the same execution count form a
# straight-line region; print the regions that carry the most warp instructions, with their first / last opcode
print("---- regions by warp instructions executed (consecutive SASS with equal execution counts)")
seen, regions = set(), []
half = [r for r in data if (r[ci["Address"]] not in seen and not seen.add(r[ci["Address"]]))]  # the CSV lists every instruction twice
cur = None
for r in half:
    ex = int(r[ci["Instructions Executed"]])
    smp = int(r[ci["# Samples"]])
    try:
        thr = float(r[ci["Avg. Threads Executed"]])
    except Exception:
        thr = 0.0
    if cur and cur["ex"] == ex:
        cur["n"] += 1; cur["smp"] += smp; cur["thr"] += thr; cur["last"] = r[ci["Source"]][:44]
        cur["ld"] += ("LDG" in r[ci["Source"]]) + ("LDL" in r[ci["Source"]]) * 0
        cur["f64"] += any(op in r[ci["Source"]] for op in ("DMUL", "DADD", "DFMA", "DSETP"))
    else:
        cur = {"ex": ex, "n": 1, "smp": smp, "thr": thr, "first": r[ci["Source"]][:44], "last": r[ci["Source"]][:44], "at": r[ci["Address"]],
               "ld": int("LDG" in r[ci["Source"]]), "f64": int(any(op in r[ci["Source"]] for op in ("DMUL", "DADD", "DFMA", "DSETP")))}
        regions.append(cur)
tot_ex = sum(g["ex"] * g["n"] for g in regions) or 1
tot_smp = sum(g["smp"] for g in regions) or 1
print("total warp instructions", tot_ex, "distinct SASS", len(half))
for g in sorted(regions, key=lambda g: -g["ex"] * g["n"])[:45]:
    print(f"{g['ex'] * g['n'] / tot_ex * 100:5.2f}% inst {g['smp'] / tot_smp * 100:5.2f}% smp  exec {g['ex']:>10} x {g['n']:>4} sass  thr {g['thr'] / g['n']:5.1f}  ldg {g['ld']:>2} f64 {g['f64']:>3}  {g['first']} .. {g['last']}")
