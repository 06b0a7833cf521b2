#!/bin/bash
# GPU session 10: streaming-softmax attention variants 24-27 (correctness on ragged shapes, timing against variant 10).
cd "$(dirname "$0")/.." || exit 1
mkdir -p gpurun_out
ST=tests/native/selftest
L=gpurun_out/s10_selftest.log
: > $L
run() { timeout 120 $ST "$@" >> $L 2>&1; rc=$?; [ $rc -ne 0 ] && echo "  -> rc=$rc for: $*" >> $L; return 0; }
for var in 24 25 26 27; do
  run attn 2 331 331 2 1 $var 0 0 5
  run attn 1 128 128 1 1 $var 0 0 1
  run attn 1 1376 1376 16 1 $var 0 0 97
  run attn 3 1376 1376 4 6 $var 0 0 61
  run attn 1 700 2900 3 8 $var 0 0 23
  run attn 1 200 70 1 1 $var 0 0 3
  run attn 1 200 17 1 1 $var 0 0 3
  run attn 1 300 113 2 1 $var 0 0 3
  run attn 1 1376 1376 16 1 $var 1 0 97
  run attn 1 3000 5000 16 2 $var 0 0 211
  run attn 2 640 9000 3 2 $var 0 0 37
done
for var in 10 24 25 26 27 12; do
  run bench_attn 1 11008 11008 16 $var
  run bench_attn 8 1376 1376 16 $var
done
for var in 10 24 25; do run bench_attn 1 22016 44032 16 $var; done
grep -E "BENCH|FAIL|rc=" $L; grep -c PASS $L
for v in 24 25; do
  timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-extras --set attn_bounded_variant=$v > gpurun_out/s10_bench_v$v.json 2> gpurun_out/s10_bench_v$v.err; echo "bench v$v rc=$?"
done
timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-extras > gpurun_out/s10_bench_v10.json 2> gpurun_out/s10_bench_v10.err
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/s10_bench_*.json')):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1])
        print(f, round(d['value'],1), 'ms', round(d['ms_per_step'],2), 'attn TF', round(d['roofline']['achieved'],1), {k:v['ms'] for k,v in d['breakdown_ms'].items() if 'attn' in k})
    except Exception as e:
        print(f, 'unreadable', e)
PY
NCU="ncu --set full --clock-control none --import-source on"
timeout 300 $ST bench_attn 1 11008 11008 16 24 > gpurun_out/ncu_plain_v24.log 2>&1 && \
timeout 600 $NCU -k regex:attn_split_kernel -s 5 -c 1 -f -o gpurun_out/s10_attn_v24 $ST bench_attn 1 11008 11008 16 24 > gpurun_out/ncu_v24.log 2>&1
echo "ncu rc=$?"
