"""Per-SOURCE-LINE view of one captured launch: where the executed instructions and the stall samples are.
usage: ncu -i X.ncu-rep --page source --csv --print-source cuda --launch-count 1 | python tools/ncu_hot_src.py [N]
(needs -lineinfo and --import-source on at capture time)"""
import csv, sys
N = int(sys.argv[1]) if len(sys.argv) > 1 else 50
rows = list(csv.reader(sys.stdin))
out = []
cur_file = ""
i = 0
while i < len(rows):
    r = rows[i]
    if r and r[0] in ("#", "Line", "Address") or (r and "Source" in r and "# Samples" in r):
        h = r
        ci = {n: k for k, n in enumerate(h)}
        i += 1
        while i < len(rows) and len(rows[i]) == len(h):
            q = rows[i]
            try:
                smp = int(q[ci["# Samples"]]); ex = int(q[ci["Instructions Executed"]])
            except Exception:
                i += 1
                continue
            thr = q[ci.get("Avg. Threads Executed", 0)] if "Avg. Threads Executed" in ci else ""
            out.append((smp, ex, thr, cur_file, q[0], q[ci["Source"]][:110]))
            i += 1
        continue
    if r and len(r) >= 2 and r[0].lower().startswith("file"):
        cur_file = r[1].split("/")[-1]
    elif r and len(r) == 1 and ("/" in r[0] or r[0].endswith((".cu", ".cuh", ".h"))):
        cur_file = r[0].split("/")[-1]
    i += 1
ts = sum(o[0] for o in out) or 1
te = sum(o[1] for o in out) or 1
print("lines", len(out), "samples", ts, "warp instructions", te)
print("---- by instructions executed")
for o in sorted(out, key=lambda o: -o[1])[:N]:
    print(f"{o[1] / te * 100:5.2f}% inst {o[0] / ts * 100:5.2f}% smp thr {o[2]:>5} {o[3]}:{o[4]:>4}  {o[5]}")
print("---- by stall samples")
for o in sorted(out, key=lambda o: -o[0])[:N]:
    print(f"{o[0] / ts * 100:5.2f}% smp {o[1] / te * 100:5.2f}% inst thr {o[2]:>5} {o[3]}:{o[4]:>4}  {o[5]}")
