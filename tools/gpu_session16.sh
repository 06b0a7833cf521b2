#!/bin/bash
# GPU session 16: shared LayerNorm of the token rows for the DPT heads (affine folded into the projections).
cd "$(dirname "$0")/.." || exit 1
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_heads_gpu.py tests/test_gs_gpu.py -m gpu -x -q > gpurun_out/s16_gpu_tests.log 2>&1; echo "pytest rc=$?"; tail -6 gpurun_out/s16_gpu_tests.log; grep -E "rel fro|saturated" gpurun_out/s16_gpu_tests.log | head -20
for sh in 1 0; do
WM_SHARE_LN=$sh timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-extras > gpurun_out/s16_bench_ln$sh.json 2> gpurun_out/s16_bench_ln$sh.err; echo "bench share_ln=$sh rc=$?"; tail -2 gpurun_out/s16_bench_ln$sh.err
done
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/s16_bench_*.json')):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1])
        print(f, round(d['value'],1), 'ms', round(d['ms_per_step'],2), 'e2e', round(d['e2e']['value'],1), d['gpu_launches'], {k:v['ms'] for k,v in d['breakdown_ms'].items() if 'dpt' in k or 'camera' in k})
    except Exception as e:
        print(f, 'unreadable', e)
PY
