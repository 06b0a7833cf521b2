#!/bin/bash
cd "$(dirname "$0")/.." || exit 1
mkdir -p gpurun_out
ST=tests/native/selftest
L=gpurun_out/s11_selftest.log
: > $L
run() { timeout 120 $ST "$@" >> $L 2>&1; rc=$?; [ $rc -ne 0 ] && echo "  -> rc=$rc for: $*" >> $L; return 0; }
for var in 24 26 27; do
  run attn 2 331 331 2 1 $var 0 0 5
  run attn 1 1376 1376 16 1 $var 0 0 97
  run attn 1 700 2900 3 8 $var 0 0 23
  run attn 1 200 70 1 1 $var 0 0 3
  run attn 1 300 113 2 1 $var 0 0 3
  run attn 1 3000 5000 16 2 $var 0 0 211
done
for var in 10 24 25 26 27; do
  run bench_attn 1 11008 11008 16 $var
  run bench_attn 8 1376 1376 16 $var
done
grep -E "BENCH|FAIL|rc=" $L; grep -c PASS $L
