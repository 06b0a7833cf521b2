#!/bin/bash
# GPU session 7: GPU tests on the final state, guard overhead and per-head streams, bench.
cd "$(dirname "$0")/.." || exit 1
mkdir -p gpurun_out
timeout 300 python tools/profile_heads.py 2>&1 | tail -1
timeout 1800 python -m pytest tests -m gpu -x -q > gpurun_out/s7_gpu_tests.log 2>&1; echo "pytest rc=$?"; tail -8 gpurun_out/s7_gpu_tests.log
for hs in 0 1; do
WM_HEAD_STREAMS=$hs timeout 900 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-extras > gpurun_out/s7_bench_hs$hs.json 2> gpurun_out/s7_bench_hs$hs.err; echo "bench head_streams=$hs rc=$?"; tail -2 gpurun_out/s7_bench_hs$hs.err
done
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/s7_bench_*.json')):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1])
        print(f, round(d['value'],1), 'ms', round(d['ms_per_step'],2), 'e2e', round(d['e2e']['value'],1), 'attn TF', round(d['roofline']['achieved'],1), {k:v['ms'] for k,v in d['breakdown_ms'].items()})
    except Exception as e:
        print(f, 'unreadable', e)
PY
