# the three render workloads, headline only (after a change to the shade classes): bash tools/r2_cls_ab.sh
for s in cornell-box veach-mis bathroom2; do
  python bench.py --scene $s --headline-only --no-cpu-baseline --steps 4 --warmup 2 > gpurun_out/cls_$s.json 2>gpurun_out/cls_$s.err
  python - $s <<'PY'
import json,sys
n=sys.argv[1]
try:
    j=json.loads(open(f'gpurun_out/cls_{n}.json').read().strip().splitlines()[-1]); r=j['roofline']
    print(f"{n:12s} spp/s {j['value']:7.1f} e2e {j['e2e']['value']:7.1f} ms {j['ms_per_step']:7.2f} ext {r['extend_ms']:6.2f} shade {r['shade_ms']:6.2f} shadow {r['shadow_ms']:6.2f} launches {j['gpu_launches']} chk {j['e2e']['image_checksum']:.9f}")
except Exception as e:
    print(n,'FAILED',e)
PY
done
