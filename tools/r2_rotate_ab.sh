# Kensler tree rotations on the SAH tree (PRT_SAH_ROTATE = passes): bash tools/r2_rotate_ab.sh
run() { # name, env, args
  env $2 python bench.py --headline-only --no-cpu-baseline --steps 3 --warmup 2 $3 > gpurun_out/rot_$1.json 2> gpurun_out/rot_$1.err
  python - "$1" <<'PY'
import json,sys
n=sys.argv[1]
try:
    j=json.loads(open(f'gpurun_out/rot_{n}.json').read().strip().splitlines()[-1]); r=j['roofline']
    print(f"{n:22s} spp/s {j['value']:7.1f} ms {j['ms_per_step']:7.2f} ext {r['extend_ms']:6.2f} shade {r['shade_ms']:6.2f} shadow {r['shadow_ms']:6.2f} nodes/ray {r['nodes_per_ray']:5.2f} tris/ray {r['tris_per_ray']:5.2f} chk {j['e2e']['image_checksum']:.9f}")
except Exception as e:
    print(n,'FAILED',e)
PY
}
for r in 0 2 8; do
  run bath_rot$r "PRT_SAH_ROTATE=$r" "--scene bathroom2"
  run cornell_rot$r "PRT_SAH_ROTATE=$r" ""
  run veach_rot$r "PRT_SAH_ROTATE=$r" "--scene veach-mis"
done
for r in 0 8; do
  echo "soup-1M rot$r"; PRT_SAH_ROTATE=$r python tools/rayquery_bench.py --tris 1 --rays 24 --mesh soup --reps 2 2>/dev/null | python -c "
import sys,json
for l in sys.stdin:
    try: j=json.loads(l); print('   ',j['case'],round(j['mrays_s']),'nodes/ray',round(j['nodes_per_ray'],2),'tris/ray',round(j['tris_per_ray'],2))
    except Exception: pass"
done
