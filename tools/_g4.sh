# A/B of ray binning (PRT_SORT_MIN) on the BVH4 build: tests, render benches, ray-query benches
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -5
export PRT_BENCH_ACCEL=bvh4
for sm in 0 262144 65536; do
export PRT_SORT_MIN=$sm
echo "== sort_min $sm"
python bench.py --no-cpu-baseline --steps 3 > gpurun_out/s4_cornell_$sm.json 2>gpurun_out/s4_err.log || tail -3 gpurun_out/s4_err.log
python bench.py --scene bathroom2 --width 1280 --height 720 --depth 50 --steps 2 --no-cpu-baseline > gpurun_out/s4_bath_$sm.json 2>gpurun_out/s4_err.log || tail -3 gpurun_out/s4_err.log
python bench.py --scene veach-mis --width 1280 --height 720 --steps 3 --no-cpu-baseline > gpurun_out/s4_veach_$sm.json 2>gpurun_out/s4_err.log || tail -3 gpurun_out/s4_err.log
if [ $sm != 65536 ]; then
python tools/rayquery_bench.py --tris 1 --rays 24 --mesh soup > gpurun_out/s4_rqs_$sm.json 2>&1
python tools/rayquery_bench.py --tris 10 --rays 24 > gpurun_out/s4_rq10_$sm.json 2>&1
fi
python - <<PY
import json
for f in ['cornell','bath','veach']:
    try:
        j=json.loads(open(f'gpurun_out/s4_{f}_$sm.json').read().strip().splitlines()[-1]); r=j['roofline']
        print(f, 'spp/s',round(j['value'],1),'e2e',round(j['e2e']['value'],1),'Mrays',round(j['mrays_s']),'ms',round(j['ms_per_step'],2),'ext/shade/shadow',[round(r[k],2) for k in ('extend_ms','shade_ms','shadow_ms')],'chk',j['e2e']['image_checksum'])
    except Exception as e: print(f,'ERR',e)
for f in ['rqs','rq10']:
    try:
        for l in open(f'gpurun_out/s4_{f}_$sm.json'):
            try:
                j=json.loads(l); print(' ',j['case'], round(j['mrays_s']), 'ms', round(j['ms'],2))
            except Exception: pass
    except Exception: pass
PY
done
