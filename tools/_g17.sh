python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 5 --warmup 3 > gpurun_out/s17_bench_n2.json 2> gpurun_out/s17_bench_n2.err; echo "n2 rc=$?"
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --steps 3 --warmup 3 --scene bathroom2 --width 1280 --height 720 --depth 50 > gpurun_out/s17_bath_n2.json 2> gpurun_out/s17_bath_n2.err; echo "bath n2 rc=$?"
tail -c 400 gpurun_out/s17_bench_n2.err
for f in gpurun_out/s17_bench_n2.json gpurun_out/s17_bath_n2.json; do python - $f <<'PY'
import json,sys
j=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1]); r=j['roofline']
print(sys.argv[1], 'n',j['n_gpus'],'spp/s',round(j['value'],1),'e2e',round(j['e2e']['value'],1),'Mrays',round(j['mrays_s']),'ms',round(j['ms_per_step'],2))
PY
done
