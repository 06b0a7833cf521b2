#!/bin/bash
# compute-sanitizer (synccheck, racecheck, memcheck) over the C++ harness on small shapes.  Every run is bounded.
cd "$(dirname "$0")/.." || exit 1
mkdir -p gpurun_out
ST=tests/native/selftest
CS=/usr/local/cuda/bin/compute-sanitizer
for tool in synccheck racecheck memcheck; do
  L=gpurun_out/s8_sanitizer_$tool.log
  : > $L
  for args in "layernorm 333" "gemm_bias 300 512 192 1 1" "gemm_resid 300 1024 192 1 0 1" "gemm_resid 600 1024 4096 1 0 7 1" "gemm_qkv 300 1 1 3" "attn 2 331 331 2 1 10 0 0 5" "attn 1 200 70 1 1 4 0 0 3" "attn 1 331 331 2 1 13 0 0 5"; do
    echo "== $tool: selftest $args" >> $L
    timeout -k 5 150 $CS --tool $tool --print-limit 20 $ST $args >> $L 2>&1
    echo "   rc=$?" >> $L
  done
  echo "---- $tool"; grep -E "ERROR SUMMARY|rc=|PASS|FAIL|hazard|Error" $L | head -40
done
