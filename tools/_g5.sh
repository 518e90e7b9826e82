# round-1 stage-4 evidence: full bench line, ncu launch list, ncu --set full of the traversal and shade kernels
cd pooraytracer_b200/csrc && true; cd ../..
python bench.py > gpurun_out/s5_bench.json 2> gpurun_out/s5_bench.err; echo "bench rc=$?"; tail -c 300 gpurun_out/s5_bench.err
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/s5_bench_ref.json 2> gpurun_out/s5_ref.err; echo "ref rc=$?"
ncu --metrics gpu__time_duration.sum --clock-control none -c 1400 --csv --log-file gpurun_out/s5_launches.csv python bench.py --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/s5_ncu.log 2>&1; echo "launchlist rc=$?"
ncu --set full --clock-control none --import-source on -k regex:k_extend -s 1 -c 3 -o gpurun_out/s5_extend -f python bench.py --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/s5_ncu2.log 2>&1; echo "extend rc=$?"
ncu --set full --clock-control none --import-source on -k regex:k_shade -s 4 -c 1 -o gpurun_out/s5_shade -f python bench.py --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/s5_ncu3.log 2>&1; echo "shade rc=$?"
ncu --set full --clock-control none --import-source on -k regex:k_shadow -s 1 -c 1 -o gpurun_out/s5_shadow -f python bench.py --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/s5_ncu4.log 2>&1; echo "shadow rc=$?"
ncu --set full --clock-control none --import-source on -k regex:k_trace_closest -s 1 -c 2 -o gpurun_out/s5_rq10 -f python tools/rayquery_bench.py --tris 10 --rays 24 --reps 1 > gpurun_out/s5_ncu5.log 2>&1; echo "rq rc=$?"
ls -la gpurun_out/*.ncu-rep
