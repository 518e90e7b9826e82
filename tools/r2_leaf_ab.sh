# leaf size / SAH triangle cost on the 1.1 M-triangle room and on cornell-box (one gpurun call): bash tools/r2_leaf_ab.sh
run() { # name, env, args
  env $2 python bench.py --headline-only --no-cpu-baseline --steps 3 --warmup 2 $3 > gpurun_out/leaf_$1.json 2> gpurun_out/leaf_$1.err
  python - "$1" <<'PY'
import json,sys
n=sys.argv[1]
try:
    j=json.loads(open(f'gpurun_out/leaf_{n}.json').read().strip().splitlines()[-1]); r=j['roofline']
    print(f"{n:22s} spp/s {j['value']:7.1f} ms {j['ms_per_step']:7.2f} ext {r['extend_ms']:6.2f} shade {r['shade_ms']:6.2f} shadow {r['shadow_ms']:6.2f} nodes/ray {r['nodes_per_ray']:5.2f} tris/ray {r['tris_per_ray']:5.2f}")
except Exception as e:
    print(n,'FAILED',e)
PY
}
for L in 0 2 3 6 8; do run bath_leaf$L "X=1" "--scene bathroom2 --max-leaf $L"; done
for C in 1.0 2.0; do run bath_ctri$C "PRT_SAH_CTRI=$C" "--scene bathroom2"; done
for L in 2 8; do run cornell_leaf$L "X=1" "--max-leaf $L"; done
