#!/bin/bash
# tools/ab_multi.sh name1 name2 ... : per variant library x {bvh2,bvh4}: cornell bench, bathroom bench, 1M soup + 10M terrain ray queries (compact lines)
mkdir -p gpurun_out
for n in "$@"; do
  lib=$PWD/pooraytracer_b200/csrc/build/variants/lib_$n.so
  [ -f "$lib" ] || { echo "missing $lib"; continue; }
  for acc in ${ACCS:-bvh2 bvh4}; do
    export PRT_BENCH_ACCEL=$acc PRT_B200_LIB=$lib
    python bench.py --no-cpu-baseline --steps 3 > gpurun_out/m_c.json 2>/dev/null
    python bench.py --scene bathroom2 --width 1280 --height 720 --depth 50 --steps 2 --no-cpu-baseline > gpurun_out/m_b.json 2>/dev/null
    python tools/rayquery_bench.py --tris 1 --rays 24 --mesh soup > gpurun_out/m_s.json 2>&1
    [ -n "$RQ10" ] && python tools/rayquery_bench.py --tris 10 --rays 24 > gpurun_out/m_t.json 2>&1
    python - "$n" "$acc" <<'PY'
import json,sys
n,acc=sys.argv[1:3]
out=f"{n:10s} {acc}"
for f,tag in (('m_c','cornell'),('m_b','bath')):
    try:
        j=json.loads(open(f'gpurun_out/{f}.json').read().strip().splitlines()[-1]); r=j['roofline']
        out+=f" | {tag} {j['value']:6.1f} spp/s {j['mrays_s']:5.0f} Mr/s ext {r['extend_ms']:5.2f} sh {r['shade_ms']:5.2f} sd {r['shadow_ms']:5.2f} chk {j['e2e']['image_checksum']:.6f}"
    except Exception as e: out+=f" | {tag} ERR"
for f in ('m_s','m_t'):
    try:
        for l in open(f'gpurun_out/{f}.json'):
            try: j=json.loads(l); out+=f" | {j['case'].split('-')[0][:4]} {j['case'].split()[1][:3]}/{j['case'].split()[2][:3]} {j['mrays_s']:5.0f}"
            except Exception: pass
    except Exception: pass
print(out)
PY
  done
done
