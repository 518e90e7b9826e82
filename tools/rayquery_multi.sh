# usage (on an N-GPU box): bash tools/rayquery_multi.sh N  -> gpurun_out/rq_nN.jsonl
# BASELINE configs[4] over N GPUs: every rank holds the mesh and traces its contiguous range of the 2^24 rays.
N=$1; : > gpurun_out/rq_n$N.jsonl
p=29600
for spec in ${SPECS:-"1 terrain contiguous" "1 soup contiguous" "10 terrain contiguous" "10 soup contiguous" "10 terrain blocks" "10 soup blocks"}; do
  set -- $spec; p=$((p+1))
  timeout 240 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $p \
    tools/rayquery_bench.py --tris $1 --rays 24 --mesh $2 --split $3 2> gpurun_out/rq_n${N}_$1_$2_$3.err | grep '^{' >> gpurun_out/rq_n$N.jsonl
  echo "$spec rc=$?"
done
python - gpurun_out/rq_n$N.jsonl <<'PY'
import json,sys
for l in open(sys.argv[1]):
    j=json.loads(l); print(j['case'], j.get('split'), 'n', j['n_gpus'], round(j['mrays_s']), 'Mrays/s  per GPU', [round(x) for x in j.get('mrays_s_per_gpu', [])])
PY
