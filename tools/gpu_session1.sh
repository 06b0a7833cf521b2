#!/bin/bash
# GPU session 1 of round 2: new attention variants (correctness + timing), MUFU rates, per-kernel ncu evidence for the
# GEMMs and the HBM-bound kernels, the same-box torch comparator, then the GPU test suite and the bench.
cd "$(dirname "$0")/.." || exit 1
mkdir -p gpurun_out
ST=tests/native/selftest
L=gpurun_out/s1_selftest.log
: > $L
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw,temperature.gpu --format=csv >> $L 2>&1
run() { timeout 180 $ST "$@" >> $L 2>&1; rc=$?; [ $rc -ne 0 ] && echo "  -> rc=$rc for: $*" >> $L; return 0; }
echo "== correctness of the pipelined variants" >> $L
for var in 13 14 15 16 17 18 19; do
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
for var in 10 11 12 13 14 15 16 17 18 19 4 7; do
  run bench_attn 1 11008 11008 16 $var
  run bench_attn 8 1376 1376 16 $var
done
for var in 10 13 15 16; do run bench_attn 1 22016 44032 16 $var; done
run bench_mufu
run bench_ln 11008
run bench_copy 11008
run bench_gather 8 518 518
run bench_gemm 1 11008 4096 1024
run bench_gemm 2 11008 3072 1024
run bench_gemm 3 11008 1024 1024
run bench_gemm 3 11008 1024 4096
grep -E "BENCH|FAIL|rc=|PASS|ok|err" $L | tail -150

echo "== torch comparator"
timeout 600 python tools/bench_torch_gpu.py --out gpurun_out/s1_torch_gpu.json 2> gpurun_out/s1_torch_gpu.err | tail -3

echo "== ncu captures (one launch each, --set full)"
NCU="ncu --set full --clock-control none --import-source on"
cap() {  # name, kernel regex, skip, selftest args...
  local name=$1 rx=$2 skip=$3; shift 3
  timeout 300 $ST "$@" > gpurun_out/ncu_plain_$name.log 2>&1 && \
  timeout 600 $NCU -k regex:$rx -s $skip -c 1 -f -o gpurun_out/s1_$name $ST "$@" > gpurun_out/ncu_$name.log 2>&1
  echo "ncu $name rc=$?"
}
cap gemm_fc1 gemm_kernel 5 bench_gemm 1 11008 4096 1024
cap gemm_qkv gemm_kernel 5 bench_gemm 2 11008 3072 1024
cap gemm_proj gemm_kernel 5 bench_gemm 3 11008 1024 1024
cap gemm_fc2 gemm_kernel 5 bench_gemm 3 11008 1024 4096
cap layernorm layernorm_kernel 5 bench_ln 11008
cap copy_cols copy_cols_kernel 5 bench_copy 11008
cap patch_gather patch_gather_kernel 5 bench_gather 8 518 518
cap attn_v10 attn_split_kernel 5 bench_attn 1 11008 11008 16 10
cap attn_v15 attn_pipe_kernel 5 bench_attn 1 11008 11008 16 15
ls -la gpurun_out/*.ncu-rep 2>/dev/null

echo "== GPU tests (default variants)"
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/s1_gpu_tests.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/s1_gpu_tests.log
echo "== bench (default) and with the pipelined attention"
timeout 900 python bench.py --steps 10 --warmup 3 > gpurun_out/s1_bench_default.json 2> gpurun_out/s1_bench_default.err; echo "bench rc=$?"
for v in 14 15 16; do
  timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --set attn_bounded_variant=$v > gpurun_out/s1_bench_v$v.json 2> gpurun_out/s1_bench_v$v.err; echo "bench v$v rc=$?"
done
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/s1_bench_*.json')):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1])
        print(f, round(d['value'],1), 'e2e', round(d['e2e']['value'],1), 'attn TF', round(d['roofline']['achieved'],1), {k:v['ms'] for k,v in d['breakdown_ms'].items()})
    except Exception as e:
        print(f, 'unreadable', e)
PY
