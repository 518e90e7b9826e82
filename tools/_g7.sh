for tp in 262144 1000000 4000000; do
export PRT_TAIL_PATHS=$tp
echo "== tail_paths $tp"
python bench.py --no-cpu-baseline --steps 3 > gpurun_out/s7_cornell_$tp.json 2>gpurun_out/s7_err.log || tail -3 gpurun_out/s7_err.log
python bench.py --scene veach-mis --width 1280 --height 720 --steps 3 --no-cpu-baseline > gpurun_out/s7_veach_$tp.json 2>gpurun_out/s7_err.log || tail -3 gpurun_out/s7_err.log
python - <<PY
import json
for f in ['cornell','veach']:
    try:
        j=json.loads(open(f'gpurun_out/s7_{f}_$tp.json').read().strip().splitlines()[-1]); r=j['roofline']
        print(f, 'spp/s',round(j['value'],1),'e2e',round(j['e2e']['value'],1),'Mrays',round(j['mrays_s']),'ms',round(j['ms_per_step'],2),'launches',j['gpu_launches'],'ext/shade/shadow',[round(r[k],2) for k in ('extend_ms','shade_ms','shadow_ms')],'chk',j['e2e']['image_checksum'])
    except Exception as e: print(f,'ERR',e)
PY
done
unset PRT_TAIL_PATHS
ncu --set full --clock-control none --import-source on -k regex:k_shade -s 1 -c 2 -o gpurun_out/s7_shade -f python bench.py --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/s7_ncu3.log 2>&1; echo "shade rc=$?"
ncu --set full --clock-control none --import-source on -k regex:k_extend -s 8 -c 2 -o gpurun_out/s7_bath_extend -f python bench.py --scene bathroom2 --width 1280 --height 720 --depth 50 --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/s7_ncu4.log 2>&1; echo "bath extend rc=$?"
