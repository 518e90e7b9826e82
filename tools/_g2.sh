mkdir -p gpurun_out
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 4 --warmup 3 > gpurun_out/s2_bench_n2.json 2> gpurun_out/s2_bench_n2.err; echo "n2 rc=$?"
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --impl reference --gpus 2 --steps 1 --warmup 0 > gpurun_out/s2_ref_n2.json 2> gpurun_out/s2_ref_n2.err; echo "ref n2 rc=$?"
python bench.py --scene veach-mis --width 1280 --height 720 --steps 3 --no-cpu-baseline > gpurun_out/s2_veach.json 2> gpurun_out/s2_veach.err; echo "veach rc=$?"
python bench.py --scene bathroom2 --width 1280 --height 720 --depth 50 --steps 3 --no-cpu-baseline > gpurun_out/s2_bath.json 2> gpurun_out/s2_bath.err; echo "bath rc=$?"
timeout 900 python tools/rayquery_bench.py --tris 10 --rays 24 > gpurun_out/s2_rq10.json 2>&1; echo "rq10 rc=$?"
tail -c 600 gpurun_out/s2_bench_n2.json; tail -3 gpurun_out/s2_bench_n2.err
