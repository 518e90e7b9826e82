# PRT_TAIL_PATHS sweep on the headline step (env knob, no rebuild): bash tools/r2_tail_ab.sh
for t in 65536 131072 262144 524288 1048576 2097152; do
  PRT_TAIL_PATHS=$t python bench.py --headline-only --no-cpu-baseline --steps 4 --warmup 2 > gpurun_out/tail_$t.json 2>/dev/null
  python - $t <<'PY'
import json,sys
t=sys.argv[1]
j=json.loads(open(f'gpurun_out/tail_{t}.json').read().strip().splitlines()[-1]); r=j['roofline']
print(f"tail_paths {int(t):8d} spp/s {j['value']:7.1f} ms {j['ms_per_step']:6.2f} ext {r['extend_ms']:6.2f} shade {r['shade_ms']:6.2f} shadow {r['shadow_ms']:6.2f} launches {j['gpu_launches']} chk {j['e2e']['image_checksum']:.9f}")
PY
done
for t in 131072 262144 1048576; do
  PRT_TAIL_PATHS=$t python bench.py --scene bathroom2 --headline-only --no-cpu-baseline --steps 3 --warmup 2 > gpurun_out/tailb_$t.json 2>/dev/null
  python - $t <<'PY'
import json,sys
t=sys.argv[1]
j=json.loads(open(f'gpurun_out/tailb_{t}.json').read().strip().splitlines()[-1]); r=j['roofline']
print(f"bathroom2 tail_paths {int(t):8d} spp/s {j['value']:7.1f} ms {j['ms_per_step']:6.2f} launches {j['gpu_launches']}")
PY
done
