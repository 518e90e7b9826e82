#!/usr/bin/env python3
"""One scene's entry of profiles/r2_traffic.json from an ncu launch list of bench.py.

  python tools/traffic_from_launches.py <launches.csv> <name it is committed under> [workload accel spp] > entry.json
  python tools/traffic_from_launches.py --merge profiles/r2_traffic.json cornell-box=entry1.json bathroom2=entry2.json ...

The list is the CSV of
  ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none
      python bench.py --steps 1 --warmup 1 --no-cpu-baseline --headline-only [--scene ...]
which renders several identical steps (warm-up, timed, stage-timed, end-to-end); the number of steps is
the number of k_tail<0, ...> launches (one per pass) and every figure is divided by it. The instrumented
step (k_*<1, ...>, the counting build of the kernels) is left out; kernels it shares with the plain steps
(k_shade, k_accumulate) are divided by one step more.
"""
import csv, json, re, sys
from collections import defaultdict

def main():
    if sys.argv[1] == "--merge":
        out = {}
        try:
            out = json.load(open(sys.argv[2]))
        except Exception:
            pass
        for kv in sys.argv[3:]:
            k, f = kv.split("=", 1)
            out[k] = json.load(open(f))
        json.dump(out, open(sys.argv[2], "w"), indent=1)
        return
    path, committed = sys.argv[1], sys.argv[2]
    workload, accel, spp = (sys.argv[3:6] + ["cornell-box", "bvh4", "64"][len(sys.argv) - 3:])[:3]
    rows = [r for r in csv.reader(l for l in open(path) if l.startswith('"'))]
    head, rows = rows[0], rows[1:]
    ik, im, iv, iid = head.index("Kernel Name"), head.index("Metric Name"), head.index("Metric Value"), head.index("ID")
    per = defaultdict(lambda: defaultdict(float))
    ids = defaultdict(set)
    for r in rows:
        m = re.search(r"prt::(k_\w+(?:<[^>]*>)?)", r[ik])
        if not m:
            continue
        k = m.group(1).replace("(bool)", "").replace("(int)", "")
        per[k][r[im]] += float(r[iv].replace(",", ""))
        ids[k].add(r[iid])
    steps = max(len(v) for k, v in ids.items() if k.startswith("k_tail<0"))
    counted = max([len(v) for k, v in ids.items() if k.startswith("k_tail<1")] + [0])
    out = {}
    for k in sorted(per):
        if re.match(r"k_(extend|shadow|tail|trace\w*|primary)<1", k):  # the counting instantiations (k_shade's first parameter is the material class)
            continue
        # kernels without the COUNT template parameter (shade, accumulate) also ran in the instrumented step
        n = steps if re.match(r"k_(extend|shadow|tail|trace|primary)", k) else steps + counted
        out[k] = {
            "launches_per_step": len(ids[k]) / n,
            "ms_per_step_under_ncu": per[k]["gpu__time_duration.sum"] / 1e6 / n,
            "dram_bytes_per_step": (per[k]["dram__bytes_read.sum"] + per[k]["dram__bytes_write.sum"]) / n,
        }
    trav = [v for k, v in out.items() if k.startswith(("k_extend", "k_shadow", "k_primary"))]
    tb = sum(v["dram_bytes_per_step"] for v in trav)
    tl = sum(v["launches_per_step"] for v in trav)
    j = {
        "what": f"dram__bytes_read.sum + dram__bytes_write.sum per kernel, summed over the launches of ONE step "
                f"({spp} spp of {workload}, {accel}), from {committed} (ncu --metrics gpu__time_duration.sum,"
                f"dram__bytes_read.sum,dram__bytes_write.sum --clock-control none python bench.py --steps 1 --warmup 1 "
                f"--no-cpu-baseline; {steps} identical steps in the list, figures divided by {steps})",
        "workload": workload, "accel": accel, "spp_per_step": int(spp),
        "traversal_dram_bytes_per_step": tb, "traversal_launches_per_step": tl,
        "traversal_dram_bytes_per_launch": tb / tl if tl else None,
        "step_dram_bytes": sum(v["dram_bytes_per_step"] for v in out.values()),
        "step_ms_under_ncu": sum(v["ms_per_step_under_ncu"] for v in out.values()),
        "source": committed,
        "kernels": out,
    }
    print(json.dumps(j, indent=1))

if __name__ == "__main__":
    main()
