# The round's evidence in one gpurun call: GPU tests, the default bench line, the ncu launch list with DRAM bytes
# (-> profiles/r1_traffic.json), one ncu --set full capture, the other two render workloads, the ray-query table.
# Outputs under gpurun_out/evidence_*; copy what is to be judged into profiles/.
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -4
python bench.py > gpurun_out/evidence_bench.json 2> gpurun_out/evidence_bench.err; echo "bench rc=$?"
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 3000 --csv --log-file gpurun_out/evidence_launches.csv python bench.py --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/evidence_ncu.log 2>&1; echo "launchlist rc=$?"
ncu --set full --clock-control none --import-source on -k regex:k_tail -s 2 -c 1 -o gpurun_out/evidence_tail -f python bench.py --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/evidence_ncu2.log 2>&1; echo "tail rc=$?"
for sc in "veach-mis --width 1280 --height 720" "bathroom2 --width 1280 --height 720 --depth 50"; do
n=$(echo $sc | cut -d' ' -f1)
python bench.py --scene $sc --steps 3 > gpurun_out/evidence_$n.json 2> gpurun_out/evidence_$n.err; echo "$n rc=$?"
done
python tools/rayquery_bench.py --tris 1 --rays 24 > gpurun_out/evidence_rq1_terrain.jsonl 2>&1
python tools/rayquery_bench.py --tris 1 --rays 24 --mesh soup > gpurun_out/evidence_rq1_soup.jsonl 2>&1
python tools/rayquery_bench.py --tris 10 --rays 24 > gpurun_out/evidence_rq10_terrain.jsonl 2>&1
python tools/rayquery_bench.py --tris 10 --rays 24 --mesh soup > gpurun_out/evidence_rq10_soup.jsonl 2>&1
for f in gpurun_out/evidence_bench.json gpurun_out/evidence_veach-mis.json gpurun_out/evidence_bathroom2.json; do python - $f <<'PY'
import json,sys
j=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1]); r=j['roofline']
print(sys.argv[1], 'spp/s',round(j['value'],1),'e2e',round(j['e2e']['value'],1),'Mrays',round(j['mrays_s']),'frac',round(r['frac'],3),'cpu',j.get('cpu_baseline',{}).get('value'))
PY
done
