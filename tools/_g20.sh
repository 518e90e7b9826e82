for ct in 0.3 0.6 1.25 2.5; do
export PRT_SAH_CTRI=$ct
for sc in "cornell-box" "bathroom2 --width 1280 --height 720 --depth 50"; do
python bench.py --scene $sc --steps 3 --no-cpu-baseline --max-leaf 8 2>/dev/null | python -c "
import sys,json
j=json.loads(sys.stdin.read().strip().splitlines()[-1]); r=j['roofline']
print('ctri $ct', j['config']['workload'][:12], 'spp/s', round(j['value'],1), 'nodes/ray', round(r['nodes_per_ray'],2), 'tris/ray', round(r['tris_per_ray'],2), 'ext/shade/shadow', [round(r[k],2) for k in ('extend_ms','shade_ms','shadow_ms')])"
done
python tools/rayquery_bench.py --tris 1 --rays 24 --mesh soup --max-leaf 8 2>/dev/null | python -c "
import sys,json
for l in sys.stdin:
    try: j=json.loads(l); print('   ',j['case'],round(j['mrays_s']), round(j['nodes_per_ray'],1), round(j['tris_per_ray'],2))
    except Exception: pass"
done
