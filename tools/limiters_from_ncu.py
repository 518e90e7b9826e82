#!/usr/bin/env python3
"""profiles/r2_limiters.json: what ncu names as the limiter of the traversal kernel, per scene, from the committed
`ncu --set full` summaries (tools/ncu_summary.py output).

  python tools/limiters_from_ncu.py cornell-box=profiles/r2_s9_extend_ncu_summary.txt bathroom2=... > profiles/r2_limiters.json

Per scene: the figures of the LAST captured launch of the file (an incoherent bounce, not bounce 0): l1tex__throughput
(% of peak), DRAM throughput (% of peak), issue slots busy, active lanes per executed instruction, and the name of the
nearest ceiling.
"""
import ast
import json
import sys


def last(line):
    return float(ast.literal_eval(line[line.index("["):])[-1])


def main():
    out = {}
    for kv in sys.argv[1:]:
        scene, path = kv.split("=", 1)
        m = {}
        for line in open(path):
            k = line.split()[0] if line.strip() else ""
            if k in ("l1tex__throughput.avg.pct_of_peak_sustained_active", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
                     "smsp__issue_active.avg.pct_of_peak_sustained_active", "smsp__thread_inst_executed_per_inst_executed.ratio",
                     "lts__throughput.avg.pct_of_peak_sustained_elapsed", "sm__warps_active.avg.pct_of_peak_sustained_active",
                     "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio", "smsp__average_warp_latency_per_inst_issued.ratio"):
                m[k] = last(line)
        l1, dram, issue = m["l1tex__throughput.avg.pct_of_peak_sustained_active"], m["gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed"], m["smsp__issue_active.avg.pct_of_peak_sustained_active"]
        ceilings = {"l1tex throughput": l1, "dram throughput": dram, "issue slots": issue, "l2 throughput": m.get("lts__throughput.avg.pct_of_peak_sustained_elapsed", 0.0)}
        top = max(ceilings, key=ceilings.get)
        ls, lat = m['smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio'], m['smsp__average_warp_latency_per_inst_issued.ratio']
        occ = m['sm__warps_active.avg.pct_of_peak_sustained_active']
        if ceilings[top] >= 70.0:
            text = (f"{top} ({ceilings[top]:.0f} % of peak) with dependent-fetch latency behind it: long-scoreboard stalls "
                    f"{ls:.1f} of {lat:.1f} cycles per issue, {occ:.0f} % warp occupancy")
        else:  # no throughput unit is near its ceiling: the walk waits on its own dependent fetches
            text = (f"dependent-fetch latency: long-scoreboard stalls {ls:.1f} of {lat:.1f} cycles per issue at {occ:.0f} % warp occupancy; "
                    f"nearest throughput ceiling {top} at {ceilings[top]:.0f} % of peak")
        out[scene] = {"limiter": text, "l1tex_pct": l1, "dram_pct": dram, "issue_active_pct": issue, "l2_pct": ceilings["l2 throughput"],
                      "simd_lanes": m["smsp__thread_inst_executed_per_inst_executed.ratio"], "warp_occupancy_pct": occ, "source": path}
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
