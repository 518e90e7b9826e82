for st in 2 3 5; do
python bench.py --scene bathroom2 --width 1280 --height 720 --depth 50 --steps $st --no-cpu-baseline > gpurun_out/s12_bath_$st.json 2>gpurun_out/s12_err.log
python - $st <<'PY'
import json,sys
j=json.loads(open(f'gpurun_out/s12_bath_{sys.argv[1]}.json').read().strip().splitlines()[-1]); r=j['roofline']
print('steps',sys.argv[1],'spp/s',round(j['value'],1),'ms',round(j['ms_per_step'],2),'e2e',round(j['e2e']['value'],1),[round(r[k],2) for k in ('extend_ms','shade_ms','shadow_ms')], j['clocks'])
PY
done
nvidia-smi --query-gpu=clocks.sm,clocks.mem,power.draw,temperature.gpu,clocks_throttle_reasons.active --format=csv
