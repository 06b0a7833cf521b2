#!/bin/bash
cd "$(dirname "$0")/.." || exit 1
mkdir -p gpurun_out
ST=tests/native/selftest
L=gpurun_out/s15_selftest.log
: > $L
run() { timeout 120 $ST "$@" >> $L 2>&1; }
for shape in "1 128 128 1" "1 256 256 16" "1 256 1376 16" "1 1376 1376 16" "2 1376 1376 16" "1 2752 1376 16" "1 1376 2752 16" "1 1376 5504 16" "8 256 1376 16" "8 1376 1376 16"; do
  run bench_attn $shape 10
done
run bench_ln 11008
run bench_copy 11008
grep BENCH $L
