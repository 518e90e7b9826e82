timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -5
for acc in bvh2 bvh4; do
export PRT_BENCH_ACCEL=$acc
echo "== $acc"
python bench.py --no-cpu-baseline --steps 3 > gpurun_out/s3_cornell_$acc.json 2>gpurun_out/s3_err.log || tail -3 gpurun_out/s3_err.log
python bench.py --scene bathroom2 --width 1280 --height 720 --depth 50 --steps 2 --no-cpu-baseline > gpurun_out/s3_bath_$acc.json 2>gpurun_out/s3_err.log || tail -3 gpurun_out/s3_err.log
python bench.py --scene veach-mis --width 1280 --height 720 --steps 3 --no-cpu-baseline > gpurun_out/s3_veach_$acc.json 2>gpurun_out/s3_err.log || tail -3 gpurun_out/s3_err.log
python tools/rayquery_bench.py --tris 1 --rays 24 --mesh soup > gpurun_out/s3_rqs_$acc.json 2>&1
python tools/rayquery_bench.py --tris 10 --rays 24 > gpurun_out/s3_rq10_$acc.json 2>&1
python - <<PY
import json
for f in ['cornell','bath','veach']:
    try:
        j=json.loads(open(f'gpurun_out/s3_{f}_$acc.json').read().strip().splitlines()[-1]); r=j['roofline']
        print(f, 'spp/s',round(j['value'],1),'Mrays',round(j['mrays_s']),'ms',round(j['ms_per_step'],2),'nodes/ray',round(r['nodes_per_ray'],1),'tris/ray',round(r['tris_per_ray'],2),'ext/shade/shadow',[round(r[k],2) for k in ('extend_ms','shade_ms','shadow_ms')],'chk',j['e2e']['image_checksum'])
    except Exception as e: print(f,'ERR',e)
for f in ['rqs','rq10']:
    for l in open(f'gpurun_out/s3_{f}_$acc.json'):
        try:
            j=json.loads(l); print(' ',j['case'], round(j['mrays_s']), 'nodes',round(j['nodes_per_ray'],1),'tris',round(j['tris_per_ray'],2))
        except Exception: pass
PY
done
