# tail kernel: tests, then A/B of PRT_TAIL_PATHS on the three render workloads
timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -5
for tp in 0 16384 49152 131072 400000; do
export PRT_TAIL_PATHS=$tp
echo "== tail_paths $tp"
python bench.py --no-cpu-baseline --steps 3 > gpurun_out/s6_cornell_$tp.json 2>gpurun_out/s6_err.log || tail -3 gpurun_out/s6_err.log
python bench.py --scene bathroom2 --width 1280 --height 720 --depth 50 --steps 2 --no-cpu-baseline > gpurun_out/s6_bath_$tp.json 2>gpurun_out/s6_err.log || tail -3 gpurun_out/s6_err.log
python bench.py --scene veach-mis --width 1280 --height 720 --steps 3 --no-cpu-baseline > gpurun_out/s6_veach_$tp.json 2>gpurun_out/s6_err.log || tail -3 gpurun_out/s6_err.log
python - <<PY
import json
for f in ['cornell','bath','veach']:
    try:
        j=json.loads(open(f'gpurun_out/s6_{f}_$tp.json').read().strip().splitlines()[-1]); r=j['roofline']
        print(f, 'spp/s',round(j['value'],1),'e2e',round(j['e2e']['value'],1),'Mrays',round(j['mrays_s']),'ms',round(j['ms_per_step'],2),'launches',j['gpu_launches'],'ext/shade/shadow',[round(r[k],2) for k in ('extend_ms','shade_ms','shadow_ms')],'chk',j['e2e']['image_checksum'])
    except Exception as e: print(f,'ERR',e)
PY
done
