# ncu --set full captures of the dominant kernels (one gpurun call, ~2 min): -> gpurun_out/s26_*.ncu-rep; condense with tools/ncu_summary.py
ncu --set full --clock-control none --import-source on -k regex:k_extend -s 25 -c 3 -o gpurun_out/s26_extend -f python bench.py --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/s26_ncu1.log 2>&1; echo "extend rc=$?"
ncu --set full --clock-control none --import-source on -k regex:k_shade -s 26 -c 2 -o gpurun_out/s26_shade -f python bench.py --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/s26_ncu2.log 2>&1; echo "shade rc=$?"
ncu --set full --clock-control none --import-source on -k regex:k_shadow -s 26 -c 1 -o gpurun_out/s26_shadow -f python bench.py --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/s26_ncu3.log 2>&1; echo "shadow rc=$?"
ncu --set full --clock-control none --import-source on -k regex:k_extend -s 14 -c 2 -o gpurun_out/s26_bath_extend -f python bench.py --scene bathroom2 --width 1280 --height 720 --depth 50 --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/s26_ncu4.log 2>&1; echo "bath rc=$?"
ncu --set full --clock-control none --import-source on -k regex:k_trace_closest -s 1 -c 2 -o gpurun_out/s26_rq10 -f python tools/rayquery_bench.py --tris 10 --rays 24 --reps 1 > gpurun_out/s26_ncu5.log 2>&1; echo "rq rc=$?"
# condense on the box: the reports themselves exceed what gpurun copies back (64 MiB)
for k in extend shade shadow bath_extend rq10; do
  ncu -i gpurun_out/s26_$k.ncu-rep --page raw --csv 2>/dev/null | python tools/ncu_summary.py > gpurun_out/s26_${k}_ncu_summary.txt
  ncu -i gpurun_out/s26_$k.ncu-rep --page source --csv --print-source sass --launch-count 1 2>/dev/null | python tools/ncu_hot.py 40 > gpurun_out/s26_${k}_hot_sass.txt
done
rm -f gpurun_out/s26_*.ncu-rep
ls -la gpurun_out | tail -15
