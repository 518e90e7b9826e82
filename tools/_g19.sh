python -m pytest tests -m gpu -x -q 2>&1 | tail -2
for ml in 2 3 4 6 8; do
for sc in "cornell-box" "bathroom2 --width 1280 --height 720 --depth 50"; do
python bench.py --scene $sc --steps 3 --no-cpu-baseline --max-leaf $ml 2>/dev/null | python -c "
import sys,json
j=json.loads(sys.stdin.read().strip().splitlines()[-1]); r=j['roofline']
print('max_leaf $ml', j['config']['workload'][:12], 'spp/s', round(j['value'],1), 'nodes/ray', round(r['nodes_per_ray'],2), 'tris/ray', round(r['tris_per_ray'],2), 'ext/shade/shadow', [round(r[k],2) for k in ('extend_ms','shade_ms','shadow_ms')])"
done
done
