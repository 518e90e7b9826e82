"""Per-SOURCE-LINE view of one captured launch: where the executed warp instructions and the stall samples are.
usage: ncu -i X.ncu-rep --page source --csv --print-source cuda,sass --launch-count 1 | python tools/ncu_hot_src.py [N]
(needs -lineinfo and --import-source on at capture time). In that CSV every source line that owns SASS appears as a
row with a line number and the metrics summed over its instructions, followed by the instruction rows themselves."""
import csv, sys
N = int(sys.argv[1]) if len(sys.argv) > 1 else 50
csv.field_size_limit(1 << 30)
out, cur_file, ci = [], "", None
for r in csv.reader(sys.stdin):
    if not r:
        continue
    if r[0] == "File Path":
        cur_file = r[1].split("/")[-1]
    elif r[0] == "Line No":
        ci = {}
        for k, n in enumerate(r):
            ci.setdefault(n, k)
    elif ci and r[0].strip().isdigit() and len(r) > ci.get("Instructions Executed", 10**9):
        try:
            smp, ex = int(r[ci["# Samples"]]), int(r[ci["Instructions Executed"]])
            thr = r[ci["Avg. Threads Executed"]]
        except Exception:
            continue
        out.append((smp, ex, thr, cur_file, r[0], r[1].strip()[:120]))
ts = sum(o[0] for o in out) or 1
te = sum(o[1] for o in out) or 1
print("source lines with code", len(out), "samples", ts, "warp instructions", te)
print("---- by warp instructions executed")
for o in sorted(out, key=lambda o: -o[1])[:N]:
    print(f"{o[1] / te * 100:5.2f}% inst {o[0] / ts * 100:5.2f}% smp thr {o[2]:>3} {o[3]}:{o[4]:>4}  {o[5]}")
print("---- by stall samples")
for o in sorted(out, key=lambda o: -o[0])[:N]:
    print(f"{o[0] / ts * 100:5.2f}% smp {o[1] / te * 100:5.2f}% inst thr {o[2]:>3} {o[3]}:{o[4]:>4}  {o[5]}")
