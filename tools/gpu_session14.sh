#!/bin/bash
# GPU session 14: cheap-items-last schedule of the whole-item attention (L = 1376 / 1374 per view).
cd "$(dirname "$0")/.." || exit 1
mkdir -p gpurun_out
ST=tests/native/selftest
L=gpurun_out/s14_selftest.log
: > $L
run() { timeout 120 $ST "$@" >> $L 2>&1; rc=$?; [ $rc -ne 0 ] && echo "  -> rc=$rc for: $*" >> $L; return 0; }
for var in 4 10 13; do
  run attn 8 1376 1376 16 1 $var 0 0 97
  run attn 8 1374 1374 16 1 $var 0 0 89
  run attn 3 1376 1376 4 6 $var 0 0 61
  run attn 2 331 331 2 1 $var 0 0 5
  run attn 5 300 113 7 1 $var 0 0 3
  run attn 1 1376 1376 16 1 $var 1 0 97
  run attn 9 640 900 20 2 $var 0 0 37
  run attn 7 128 128 30 1 $var 0 0 1
  run attn 6 384 500 31 1 $var 0 0 7
done
for var in 10 4; do
  run bench_attn 8 1376 1376 16 $var
  run bench_attn 8 1374 1374 16 $var
  run bench_attn 1 11008 11008 16 $var
  run bench_attn 4 1376 1376 16 $var
  run bench_attn 16 1376 1376 16 $var
done
grep -E "BENCH|FAIL|rc=" $L; grep -c PASS $L
timeout 900 python -m pytest tests/test_parity_gpu.py -m gpu -x -q > gpurun_out/s14_gpu_tests.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/s14_gpu_tests.log
timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-extras > gpurun_out/s14_bench.json 2> gpurun_out/s14_bench.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.loads(open('gpurun_out/s14_bench.json').read().strip().splitlines()[-1])
print(round(d['value'],1), 'ms', round(d['ms_per_step'],2), 'e2e', round(d['e2e']['value'],1), 'attn TF', round(d['roofline']['achieved'],1), {k:v['ms'] for k,v in d['breakdown_ms'].items()})
PY
