timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -4
for acc in lbvh; do
echo "== $acc"
python bench.py --no-cpu-baseline --steps 3 --accel $acc > gpurun_out/s9_cornell_$acc.json 2>gpurun_out/s9_err.log || tail -3 gpurun_out/s9_err.log
python bench.py --scene bathroom2 --width 1280 --height 720 --depth 50 --steps 2 --no-cpu-baseline --accel $acc > gpurun_out/s9_bath_$acc.json 2>gpurun_out/s9_err.log || tail -3 gpurun_out/s9_err.log
python tools/rayquery_bench.py --tris 10 --rays 24 --accel $acc > gpurun_out/s9_rq10_$acc.json 2>&1
python tools/rayquery_bench.py --tris 1 --rays 24 --mesh soup --accel $acc > gpurun_out/s9_rqs_$acc.json 2>&1
python - <<PY
import json
for f in ['cornell','bath']:
    try:
        j=json.loads(open(f'gpurun_out/s9_{f}_$acc.json').read().strip().splitlines()[-1]); r=j['roofline']
        print(f, 'spp/s',round(j['value'],1),'Mrays',round(j['mrays_s']),'ms',round(j['ms_per_step'],2),'nodes/ray',round(r['nodes_per_ray'],1),'tris/ray',round(r['tris_per_ray'],2),'create/commit s',round(j['config']['scene_create_s'],3),round(j['config']['scene_commit_s'],3),'chk',j['e2e']['image_checksum'])
    except Exception as e: print(f,'ERR',e)
for f in ['rq10','rqs']:
    for l in open(f'gpurun_out/s9_{f}_$acc.json'):
        try:
            j=json.loads(l); print(' ',j['case'], round(j['mrays_s']), 'nodes',round(j['nodes_per_ray'],1),'tris',round(j['tris_per_ray'],2),'create/commit s',round(j['scene_create_s'],2),round(j['scene_commit_s'],2), 'n_nodes', j['n_nodes'], 'build_ms', j.get('build_ms'))
        except Exception: pass
PY
done
