python -m pytest tests -m gpu -x -q 2>&1 | tail -3
for acc in lbvh; do
echo "== $acc"
for sc in "cornell-box" "bathroom2 --width 1280 --height 720 --depth 50"; do
python bench.py --scene $sc --steps 3 --no-cpu-baseline --accel $acc 2>/dev/null | python -c "
import sys,json
j=json.loads(sys.stdin.read().strip().splitlines()[-1]); r=j['roofline']
print(j['config']['workload'][:12], 'spp/s', round(j['value'],1), 'nodes/ray', round(r['nodes_per_ray'],2), 'tris/ray', round(r['tris_per_ray'],2), 'commit_s', round(j['config']['scene_commit_s'],3), 'chk', j['e2e']['image_checksum'])"
done
for m in "--tris 1 --mesh soup" "--tris 10"; do
python tools/rayquery_bench.py --rays 24 $m --accel $acc 2>/dev/null | python -c "
import sys,json
for l in sys.stdin:
    try: j=json.loads(l); print('   ',j['case'],round(j['mrays_s']), round(j['nodes_per_ray'],1), round(j['tris_per_ray'],2), 'commit_s', round(j['scene_commit_s'],2), 'build_ms', round(j['build_ms'],2))
    except Exception: pass"
done
done
