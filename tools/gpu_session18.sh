#!/bin/bash
cd "$(dirname "$0")/.." || exit 1
mkdir -p gpurun_out
ST=tests/native/selftest
L=gpurun_out/s18_selftest.log
: > $L
run() { timeout 120 $ST "$@" >> $L 2>&1; rc=$?; [ $rc -ne 0 ] && echo "  -> rc=$rc for: $*" >> $L; return 0; }
for var in 4 10 7 12; do
  run attn 2 331 331 2 1 $var 0 0 5
  run attn 1 1376 1376 16 1 $var 0 0 97
  run attn 1 700 2900 3 8 $var 0 0 23
  run attn 1 200 70 1 1 $var 0 0 3
  run attn 5 300 113 7 1 $var 0 0 3
  run attn 1 3000 5000 16 2 $var 0 0 211
  run attn 2 640 9000 3 2 $var 0 0 37
  run attn 8 1374 1374 16 1 $var 0 0 89
done
for var in 10 4 12; do
  run bench_attn 1 11008 11008 16 $var
  run bench_attn 8 1376 1376 16 $var
  run bench_attn 1 22016 44032 16 $var
done
grep -E "BENCH|FAIL|rc=" $L; grep -c PASS $L
timeout 900 python -m pytest tests/test_parity_gpu.py -m gpu -x -q > gpurun_out/s18_gpu_tests.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/s18_gpu_tests.log
timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-extras > gpurun_out/s18_bench.json 2> gpurun_out/s18_bench.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.loads(open('gpurun_out/s18_bench.json').read().strip().splitlines()[-1])
print(round(d['value'],1), 'ms', round(d['ms_per_step'],2), 'e2e', round(d['e2e']['value'],1), 'attn TF', round(d['roofline']['achieved'],1), {k:v['ms'] for k,v in d['breakdown_ms'].items()})
PY
