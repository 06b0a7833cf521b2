#!/bin/bash
# Stream-K schedule of the residual GEMMs: correctness + determinism (native selftest), then isolated timings with the
# schedule on / off, then the bench line.
cd "$(dirname "$0")/.." || exit 1
mkdir -p gpurun_out
S=tests/native/selftest
L=gpurun_out/s30_streamk.log
: > $L
for args in "11008 1024 1024 1 0 37 1" "11008 1024 4096 1 0 37 1" "5504 1024 1024 1 0 37 1" "11008 1024 1024 0 0 37 1" \
            "22016 1024 1024 1 0 61 1" "10992 1024 4096 1 0 37 1" "1000 1024 2048 1 0 7 1" "19264 1024 1024 1 0 61 1"; do
  timeout 300 $S gemm_resid $args >> $L 2>&1; echo "rc=$? ($args)" >> $L
done
for sk in 0 1; do
  echo "WM_GEMM_STREAMK=$sk" >> $L
  for shape in "11008 1024 1024" "11008 1024 4096" "5504 1024 1024" "5504 1024 4096" "22016 1024 1024" "22016 1024 4096"; do
    WM_GEMM_STREAMK=$sk timeout 120 $S bench_gemm 4 $shape >> $L 2>&1
  done
done
timeout 120 $S bench_gemm 3 11008 1024 1024 >> $L 2>&1
timeout 120 $S bench_gemm 3 11008 1024 4096 >> $L 2>&1
cat $L
timeout 600 python bench.py --no-cpu-baseline --no-extras > gpurun_out/s30_bench_streamk.json 2> gpurun_out/s30_bench.err; echo "bench rc=$?"
WM_GEMM_STREAMK=0 timeout 600 python bench.py --no-cpu-baseline --no-extras > gpurun_out/s30_bench_nostreamk.json 2>> gpurun_out/s30_bench.err; echo "bench rc=$?"
timeout 900 python -m pytest tests/test_parity_gpu.py -m gpu -x -q > gpurun_out/s30_parity_tests.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/s30_parity_tests.log
