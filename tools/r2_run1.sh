#!/bin/bash
# round 2, GPU call 1: full -m gpu suite on the new record layout / primary-hit reuse / smem stack, then A/B
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv,noheader > gpurun_out/r2_1_smi.txt
( time timeout 1500 python -m pytest tests -m gpu -x -q --durations=15 ) > gpurun_out/r2_1_pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r2_1_pytest.log
tail -5 gpurun_out/r2_1_pytest.log
timeout 300 python bench.py --no-cpu-baseline --steps 5 --warmup 3 > gpurun_out/r2_1_bench_default.json 2> gpurun_out/r2_1_bench_default.err
PRT_PRIMARY_REUSE=0 timeout 300 python bench.py --no-cpu-baseline --steps 5 --warmup 3 > gpurun_out/r2_1_bench_noreuse.json 2> gpurun_out/r2_1_bench_noreuse.err
python - <<'PY'
import json
for n in ("default","noreuse"):
    try:
        j=json.loads(open(f'gpurun_out/r2_1_bench_{n}.json').read().strip().splitlines()[-1]); r=j['roofline']
        print(f"{n:10s} spp/s {j['value']:7.1f} e2e {j['e2e']['value']:7.1f} Mrays/s {j['mrays_s']:7.0f} ms {j['ms_per_step']:6.2f} ext {r['extend_ms']:6.2f} shade {r['shade_ms']:6.2f} shadow {r['shadow_ms']:6.2f} chk {j['e2e']['image_checksum']:.9f}")
    except Exception as e:
        print(n,"FAILED",e)
PY
AB_BATH=1 bash tools/ab_bench.sh rq "$@" 2>&1 | tee gpurun_out/r2_1_ab.txt
