# ncu --set full captures of the dominant kernels (one gpurun call, ~3 min): condensed on the box with tools/ncu_summary.py
# and tools/ncu_hot.py into gpurun_out/${TAG}_*; usage: TAG=r2_s2 bash tools/ncu_captures.sh
# launches per 64-spp cornell step: k_primary 1, k_extend 24 (bounces 1..24), k_shade 25, k_shadow 25, k_tail 1 (bounces 25..100);
# bench.py runs an instrumented step first (the COUNT instantiations match the same regex), so skip one step's worth.
TAG=${TAG:-r2}
B="python bench.py --steps 1 --warmup 1 --no-cpu-baseline --headline-only"
N="ncu --set full --clock-control none --import-source on"
$N -k regex:k_extend -s 24 -c 2 -o gpurun_out/${TAG}_extend -f $B > gpurun_out/${TAG}_ncu1.log 2>&1; echo "extend rc=$?"
$N -k regex:k_shade -s 25 -c 2 -o gpurun_out/${TAG}_shade -f $B > gpurun_out/${TAG}_ncu2.log 2>&1; echo "shade rc=$?"
$N -k regex:k_shadow -s 25 -c 2 -o gpurun_out/${TAG}_shadow -f $B > gpurun_out/${TAG}_ncu3.log 2>&1; echo "shadow rc=$?"
if [ -z "$NCU_SKIP_BATH" ]; then
$N -k regex:k_extend -s 13 -c 2 -o gpurun_out/${TAG}_bath_extend -f python bench.py --scene bathroom2 --steps 1 --warmup 1 --no-cpu-baseline --headline-only > gpurun_out/${TAG}_ncu4.log 2>&1; echo "bath rc=$?"
$N -k regex:k_trace_closest -s 1 -c 2 -o gpurun_out/${TAG}_rq10 -f python tools/rayquery_bench.py --tris 10 --rays 24 --reps 1 > gpurun_out/${TAG}_ncu5.log 2>&1; echo "rq rc=$?"
fi
# condense on the box: the reports themselves exceed what gpurun copies back (64 MiB)
for k in extend shade shadow bath_extend rq10; do
  [ -f gpurun_out/${TAG}_$k.ncu-rep ] || continue
  ncu -i gpurun_out/${TAG}_$k.ncu-rep --page raw --csv 2>/dev/null | python tools/ncu_summary.py > gpurun_out/${TAG}_${k}_ncu_summary.txt
  ncu -i gpurun_out/${TAG}_$k.ncu-rep --page source --csv --print-source sass --launch-count 1 2>/dev/null | python tools/ncu_hot.py 60 > gpurun_out/${TAG}_${k}_hot_sass.txt
  ncu -i gpurun_out/${TAG}_$k.ncu-rep --page source --csv --print-source cuda,sass --launch-count 1 2>/dev/null > gpurun_out/${TAG}_${k}_src_raw.csv
  python tools/ncu_hot_src.py 60 < gpurun_out/${TAG}_${k}_src_raw.csv > gpurun_out/${TAG}_${k}_hot_src.txt 2>&1
  rm -f gpurun_out/${TAG}_${k}_src_raw.csv
done
rm -f gpurun_out/${TAG}_*.ncu-rep
ls -la gpurun_out | tail -15
