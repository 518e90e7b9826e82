#!/bin/bash
# A/B harness: runs bench.py (and optionally the ray-query microbench) once per variant library in
# pooraytracer_b200/csrc/build/variants/ (same ABI, different -D flags), via PRT_B200_LIB.
#   tools/ab_bench.sh [rq] name1 name2 ...     -> gpurun_out/ab_<name>.json
mkdir -p gpurun_out
RQ=0; if [ "$1" = "rq" ]; then RQ=1; shift; fi
for n in "$@"; do
  lib=$PWD/pooraytracer_b200/csrc/build/variants/lib_$n.so
  [ -f "$lib" ] || { echo "missing $lib"; continue; }
  PRT_B200_LIB=$lib timeout 300 python bench.py --no-cpu-baseline --headline-only --steps 4 --warmup 3 > gpurun_out/ab_$n.json 2> gpurun_out/ab_$n.err
  python - "$n" <<'PY'
import json,sys
n=sys.argv[1]
try:
    j=json.loads(open(f'gpurun_out/ab_{n}.json').read().strip().splitlines()[-1]); r=j['roofline']
    print(f"{n:14s} spp/s {j['value']:7.1f} e2e {j['e2e']['value']:7.1f} Mrays/s {j['mrays_s']:7.0f} ms {j['ms_per_step']:6.2f} ext {r['extend_ms']:6.2f} shade {r['shade_ms']:6.2f} shadow {r['shadow_ms']:6.2f} frac {r['frac']:.3f} chk {j['e2e']['image_checksum']:.9f}")
except Exception as e:
    print(n,'FAILED',e)
PY
  if [ -n "$AB_BATH" ]; then
    PRT_B200_LIB=$lib timeout 300 python bench.py --scene bathroom2 --width 1280 --height 720 --depth 50 --steps ${AB_BATH_STEPS:-2} --no-cpu-baseline --headline-only > gpurun_out/ab_bath_$n.json 2> gpurun_out/ab_bath_$n.err
    python - "$n" <<'PY'
import json,sys
n=sys.argv[1]
try:
    j=json.loads(open(f'gpurun_out/ab_bath_{n}.json').read().strip().splitlines()[-1]); r=j['roofline']
    print(f"   bathroom2   spp/s {j['value']:7.1f} Mrays/s {j['mrays_s']:7.0f} ms {j['ms_per_step']:6.2f} ext {r['extend_ms']:6.2f} shade {r['shade_ms']:6.2f} shadow {r['shadow_ms']:6.2f} chk {j['e2e']['image_checksum']:.9f}")
except Exception as e:
    print(n,'bath FAILED',e)
PY
  fi
  if [ $RQ = 1 ]; then
    PRT_B200_LIB=$lib timeout 300 python tools/rayquery_bench.py --tris 1 --rays 24 --mesh soup 2>/dev/null | python -c "
import sys,json
for l in sys.stdin:
    try: j=json.loads(l); print('   ',j['case'],round(j['mrays_s']))
    except Exception: pass"
  fi
done
