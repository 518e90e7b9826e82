# The round's evidence in one gpurun call (TAG=r2_sN bash tools/evidence_run.sh): GPU tests, the default bench line, the ncu launch
# lists with DRAM bytes of the three render workloads (-> profiles/r2_traffic.json via tools/traffic_from_launches.py), and the
# ncu --set full captures (tools/ncu_captures.sh -> profiles/r2_limiters.json via tools/limiters_from_ncu.py).
# Outputs under gpurun_out/${TAG}_*; copy what is to be judged into profiles/.
TAG=${TAG:-r2_ev}
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
( time python bench.py > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err ) 2>&1 | grep real; echo "bench rc=$?"
M="ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 3000 --csv"
$M --log-file gpurun_out/${TAG}_launches_cornell-box.csv python bench.py --steps 1 --warmup 1 --no-cpu-baseline --headline-only > gpurun_out/${TAG}_ncu_l1.log 2>&1; echo "launchlist cornell rc=$?"
$M --log-file gpurun_out/${TAG}_launches_veach-mis.csv python bench.py --scene veach-mis --steps 1 --warmup 1 --no-cpu-baseline --headline-only > gpurun_out/${TAG}_ncu_l2.log 2>&1; echo "launchlist veach rc=$?"
$M --log-file gpurun_out/${TAG}_launches_bathroom2.csv python bench.py --scene bathroom2 --steps 1 --warmup 1 --no-cpu-baseline --headline-only > gpurun_out/${TAG}_ncu_l3.log 2>&1; echo "launchlist bathroom rc=$?"
for s in cornell-box veach-mis bathroom2; do
  python tools/traffic_from_launches.py gpurun_out/${TAG}_launches_$s.csv profiles/${TAG}_launches_$s.csv $s auto 64 > gpurun_out/${TAG}_traffic_$s.json
done
TAG=$TAG bash tools/ncu_captures.sh
N="ncu --set full --clock-control none --import-source on"
$N -k regex:k_extend -s 24 -c 2 -o gpurun_out/${TAG}_veach_extend -f python bench.py --scene veach-mis --steps 1 --warmup 1 --no-cpu-baseline --headline-only > gpurun_out/${TAG}_ncu6.log 2>&1; echo "veach extend rc=$?"
ncu -i gpurun_out/${TAG}_veach_extend.ncu-rep --page raw --csv 2>/dev/null | python tools/ncu_summary.py > gpurun_out/${TAG}_veach_extend_ncu_summary.txt
rm -f gpurun_out/${TAG}_*.ncu-rep
ls gpurun_out | grep ${TAG} | head -60
