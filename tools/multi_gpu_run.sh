# usage (on an N-GPU box): bash tools/multi_gpu_run.sh N  -> gpurun_out/s25_*_nN.json (cornell-box and bathroom2 through torchrun)
N=$1
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/s25_bench_n$N.json 2> gpurun_out/s25_bench_n$N.err; echo "n$N rc=$?"
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N --steps 3 --warmup 3 --no-cpu-baseline --scene bathroom2 --width 1280 --height 720 --depth 50 > gpurun_out/s25_bath_n$N.json 2> gpurun_out/s25_bath_n$N.err; echo "bath n$N rc=$?"
for f in gpurun_out/s25_bench_n$N.json gpurun_out/s25_bath_n$N.json; do python - $f <<'PY'
import json,sys
j=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
print(sys.argv[1], 'n',j['n_gpus'],'spp/s',round(j['value'],1),'e2e',round(j['e2e']['value'],1),'Mrays',round(j['mrays_s']),'ms',round(j['ms_per_step'],2), j['roofline']['traffic'])
PY
done
