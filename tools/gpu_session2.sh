#!/bin/bash
# GPU session 2 of round 2: the two-issuer attention kernel (variants 13-23) against variants 10 / 4, the GEMM tail
# split (correctness, determinism, timing), the GPU test suite, the bench with the candidates, ncu of the candidates.
cd "$(dirname "$0")/.." || exit 1
mkdir -p gpurun_out
ST=tests/native/selftest
L=gpurun_out/s2_selftest.log
: > $L
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw,temperature.gpu --format=csv >> $L 2>&1
run() { timeout 180 $ST "$@" >> $L 2>&1; rc=$?; [ $rc -ne 0 ] && echo "  -> rc=$rc for: $*" >> $L; return 0; }
echo "== correctness of the pipelined variants" >> $L
for var in 13 14 15 16 17 18 19 20 21 22 23; do
  run attn 2 331 331 2 1 $var 0 0 5
  run attn 1 128 128 1 1 $var 0 0 1
  run attn 1 1376 1376 16 1 $var 0 0 97
  run attn 3 1376 1376 4 6 $var 0 0 61
  run attn 1 700 2900 3 8 $var 0 0 23
  run attn 1 200 70 1 1 $var 0 0 3
  run attn 1 1376 1376 16 1 $var 1 0 97
  run attn 1 3000 5000 16 2 $var 0 0 211
  run attn 2 640 9000 3 2 $var 0 0 37
done
echo "== timing" >> $L
for var in 10 13 14 15 16 17 18 19 4 20 21 22 23; do
  run bench_attn 1 11008 11008 16 $var
  run bench_attn 8 1376 1376 16 $var
done
for var in 10 13 15 17; do run bench_attn 1 22016 44032 16 $var; done
echo "== GEMM tail split" >> $L
run gemm_resid 11008 1024 1024 1 0 37 1
run gemm_resid 11008 1024 4096 1 0 37 1
run gemm_resid 300 1024 192 1 0 1 1
run gemm_resid 5000 1024 1024 0 0 13 1
run gemm_resid 22016 1024 4096 1 0 97 1
run gemm_resid 1370 1024 1024 1 0 7 1
run bench_gemm 3 11008 1024 1024
run bench_gemm 4 11008 1024 1024
run bench_gemm 3 11008 1024 4096
run bench_gemm 4 11008 1024 4096
run bench_gemm 3 5504 1024 4096
run bench_gemm 4 5504 1024 4096
grep -E "BENCH|FAIL|rc=|PASS|ok|err" $L | tail -170

echo "== GPU tests (default variants)"
timeout 1800 python -m pytest tests -m gpu -x -q > gpurun_out/s2_gpu_tests.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/s2_gpu_tests.log
echo "== bench: default, GEMM split, attention candidates"
timeout 900 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/s2_bench_default.json 2> gpurun_out/s2_bench_default.err; echo "bench rc=$?"
timeout 900 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --set gemm_split=1 > gpurun_out/s2_bench_split.json 2> gpurun_out/s2_bench_split.err; echo "bench split rc=$?"
for v in 13 15 17; do
  timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --set attn_bounded_variant=$v --set gemm_split=1 > gpurun_out/s2_bench_v$v.json 2> gpurun_out/s2_bench_v$v.err; echo "bench v$v rc=$?"
done
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/s2_bench_*.json')):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1])
        print(f, round(d['value'],1), 'e2e', round(d['e2e']['value'],1), 'attn TF', round(d['roofline']['achieved'],1), {k:v['ms'] for k,v in d['breakdown_ms'].items()})
    except Exception as e:
        print(f, 'unreadable', e)
PY

echo "== ncu captures (one launch each, --set full)"
NCU="ncu --set full --clock-control none --import-source on"
cap() {  # name, kernel regex, skip, selftest args...
  local name=$1 rx=$2 skip=$3; shift 3
  timeout 300 $ST "$@" > gpurun_out/ncu_plain_$name.log 2>&1 && \
  timeout 600 $NCU -k regex:$rx -s $skip -c 1 -f -o gpurun_out/s2_$name $ST "$@" > gpurun_out/ncu_$name.log 2>&1
  echo "ncu $name rc=$?"
}
cap attn_v13 attn_pipe_kernel 5 bench_attn 1 11008 11008 16 13
cap attn_v17 attn_pipe_kernel 5 bench_attn 1 11008 11008 16 17
cap attn_v22_frame attn_pipe_kernel 5 bench_attn 8 1376 1376 16 22
cap gemm_proj_split gemm_kernel 5 bench_gemm 4 11008 1024 1024
ls -la gpurun_out/*.ncu-rep 2>/dev/null
