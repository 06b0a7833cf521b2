#!/bin/bash
# Runs the native self-test matrix on the GPU box; one process per case, each under its own timeout.
# Usage: tests/native/run_selftest.sh [probe|correct|bench|all]   (log: gpurun_out/selftest.log)
cd "$(dirname "$0")/../.." || exit 1
mkdir -p gpurun_out
LOG=gpurun_out/selftest.log
ST=tests/native/selftest
what=${1:-all}
fails=0
run() {
  timeout 180 $ST "$@" >> $LOG 2>&1
  rc=$?
  if [ $rc -ne 0 ]; then echo "  -> rc=$rc for: $*" >> $LOG; fails=$((fails+1)); fi
}
echo "== selftest $(date -u) $(nvidia-smi --query-gpu=name --format=csv,noheader | head -1)" >> $LOG
if [ "$what" = probe ] || [ "$what" = all ]; then
  for mode in 0 1; do run probe $mode 1024 1024; done
fi
if [ "$what" = correct ] || [ "$what" = all ]; then
  run layernorm 333
  run gemm_bias 300 512 192 0 1
  run gemm_bias 300 512 192 1 1
  run gemm_bias 5137 1024 256 0 37
  run gemm_bias 700 256 4096 1 7
  run gemm_resid 300 512 192 1 0 1
  run gemm_resid 300 1024 192 0 1 1
  run gemm_resid 5137 1024 1024 1 1 41
  run gemm_table 300 1024 588 640 77
  run gemm_qkv 300 1 1 3
  run gemm_qkv 300 0 0 3
  run gemm_qkv 2500 1 1 29
  for var in 0 1 2 3; do
    run attn 2 331 331 2 1 $var 0 0 5
    run attn 1 128 128 1 1 $var 0 0 1
    run attn 1 1376 1376 16 1 $var 0 0 97
    run attn 3 1376 1376 4 6 $var 0 0 61
    run attn 1 512 11008 2 1 $var 0 0 31
    run attn 1 700 2900 3 8 $var 0 0 23
  done
fi
if [ "$what" = attn ]; then
  for var in ${VARIANTS:-4 7 8 9 10 11 12}; do
    run attn 2 331 331 2 1 $var 0 0 5
    run attn 1 128 128 1 1 $var 0 0 1
    run attn 1 1376 1376 16 1 $var 0 0 97
    run attn 3 1376 1376 4 6 $var 0 0 61
    run attn 1 512 11008 2 1 $var 0 0 31
    run attn 1 700 2900 3 8 $var 0 0 23
    run attn 1 200 70 1 1 $var 0 0 3
    run attn 1 1376 1376 16 1 $var 1 0 97
    run attn 1 3000 5000 16 2 $var 0 0 211
    run attn 2 640 9000 3 2 $var 0 0 37
  done
  for var in ${VARIANTS:-4 7 8 9 10 11 12} 2; do
    run bench_attn 8 1376 1376 16 $var
    run bench_attn 1 11008 11008 16 $var
    run bench_attn 1 22016 44032 16 $var
    run bench_attn 8 1376 1376 16 $var 1
    run bench_attn 1 11008 11008 16 $var 1
  done
fi
if [ "$what" = bench ] || [ "$what" = all ]; then
  run bench_ln 11008
  run bench_gemm 0 11008 1024 1024
  run bench_gemm 1 11008 4096 1024
  run bench_gemm 2 11008 3072 1024
  run bench_gemm 3 11008 1024 1024
  run bench_gemm 3 11008 1024 4096
  run bench_gemm 0 11008 4096 4096
  for var in 0 1 2 3; do
    run bench_attn 8 1376 1376 16 $var
    run bench_attn 1 11008 11008 16 $var
    run bench_attn 1 22016 44032 16 $var
  done
fi
echo "== done, failing cases: $fails" >> $LOG
tail -n 120 $LOG
exit 0
