# accel A/B on the render workloads: spp/s, stage times, nodes / triangles per ray
mkdir -p gpurun_out
for acc in bvh4 bvh4q bvh2; do
  for sc in cornell-box bathroom2 veach-mis; do
    python bench.py --accel $acc --scene $sc --no-cpu-baseline --headline-only --steps 3 --warmup 2 > gpurun_out/acc_${acc}_${sc}.json 2> gpurun_out/acc_${acc}_${sc}.err
    python - $acc $sc <<'PY'
import json,sys
a,s=sys.argv[1:3]
try:
    j=json.loads(open(f'gpurun_out/acc_{a}_{s}.json').read().strip().splitlines()[-1]); r=j['roofline']
    print(f"{a:6s} {s:12s} spp/s {j['value']:7.1f} ms {j['ms_per_step']:7.2f} ext {r['extend_ms']:7.2f} shade {r['shade_ms']:6.2f} shadow {r['shadow_ms']:6.2f} nodes/ray {r['nodes_per_ray']:6.2f} tris/ray {r['tris_per_ray']:5.2f} B/ray {r['bytes_per_ray']:7.1f}")
except Exception as e:
    print(a,s,'FAILED',e)
PY
  done
done
