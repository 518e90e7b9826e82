# SAH builder variants (bins per axis, full sweep for small subtrees) on bathroom2 / cornell-box / soup-1M ray queries: bash tools/r2_builder_ab.sh
run() { # name, env, args
  env $2 python bench.py --headline-only --no-cpu-baseline --steps 3 --warmup 2 $3 > gpurun_out/bld_$1.json 2> gpurun_out/bld_$1.err
  python - "$1" <<'PY'
import json,sys
n=sys.argv[1]
try:
    j=json.loads(open(f'gpurun_out/bld_{n}.json').read().strip().splitlines()[-1]); r=j['roofline']
    print(f"{n:26s} spp/s {j['value']:7.1f} ms {j['ms_per_step']:7.2f} ext {r['extend_ms']:6.2f} shade {r['shade_ms']:6.2f} shadow {r['shadow_ms']:6.2f} nodes/ray {r['nodes_per_ray']:5.2f} tris/ray {r['tris_per_ray']:5.2f} commit {j['config']['scene_commit_s']:.2f} s")
except Exception as e:
    print(n,'FAILED',e)
PY
}
for v in "base:X=1" "bins32:PRT_SAH_BINS=32" "bins64:PRT_SAH_BINS=64" "sweep64:PRT_SAH_SWEEP_BELOW=64" "sweep1k:PRT_SAH_SWEEP_BELOW=1024" "bins32sweep1k:PRT_SAH_BINS=32 PRT_SAH_SWEEP_BELOW=1024" "bins64sweep4k:PRT_SAH_BINS=64 PRT_SAH_SWEEP_BELOW=4096"; do
  n=${v%%:*}; e=${v#*:}
  run bath_$n "$e" "--scene bathroom2"
done
for v in "base:X=1" "bins32sweep1k:PRT_SAH_BINS=32 PRT_SAH_SWEEP_BELOW=1024" "bins64sweep4k:PRT_SAH_BINS=64 PRT_SAH_SWEEP_BELOW=4096"; do
  n=${v%%:*}; e=${v#*:}
  run cornell_$n "$e" ""
  run veach_$n "$e" "--scene veach-mis"
  echo "soup-1M $n"; env $e python tools/rayquery_bench.py --tris 1 --rays 24 --mesh soup --reps 2 2>/dev/null | python -c "
import sys,json
for l in sys.stdin:
    try: j=json.loads(l); print('   ',j['case'],round(j['mrays_s']),'nodes/ray',round(j['nodes_per_ray'],2),'tris/ray',round(j['tris_per_ray'],2))
    except Exception: pass"
done
