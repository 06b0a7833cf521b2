#!/bin/bash
# Two staging tiles per epilogue team in the GEMMs: native correctness cases, isolated timings, parity tests, bench line.
cd "$(dirname "$0")/.." || exit 1
mkdir -p gpurun_out
S=tests/native/selftest
L=gpurun_out/s36_gemm_stage.log
: > $L
timeout 120 $S gemm_bias 300 512 192 0 1 >> $L 2>&1; echo "rc=$?" >> $L
timeout 120 $S gemm_bias 1000 4096 1024 1 1 >> $L 2>&1; echo "rc=$?" >> $L
timeout 120 $S gemm_resid 1000 1024 2048 1 0 7 1 >> $L 2>&1; echo "rc=$?" >> $L
timeout 120 $S gemm_resid 11008 1024 4096 1 0 61 1 >> $L 2>&1; echo "rc=$?" >> $L
timeout 120 $S gemm_resid 11008 1024 1024 1 1 61 0 >> $L 2>&1; echo "rc=$?" >> $L
timeout 120 $S gemm_table >> $L 2>&1; echo "rc=$?" >> $L
timeout 120 $S gemm_qkv 300 1 1 3 >> $L 2>&1; echo "rc=$?" >> $L
timeout 120 $S gemm_qkv 2752 1 1 5 >> $L 2>&1; echo "rc=$?" >> $L
for rep in 1 2; do
timeout 60 $S bench_gemm 2 11008 3072 1024 >> $L 2>&1
timeout 60 $S bench_gemm 1 11008 4096 1024 >> $L 2>&1
timeout 60 $S bench_gemm 3 11008 1024 1024 >> $L 2>&1
timeout 60 $S bench_gemm 4 11008 1024 4096 >> $L 2>&1
timeout 60 $S bench_gemm 0 11008 4096 4096 >> $L 2>&1
done
cat $L
timeout 900 python -m pytest tests/test_parity_gpu.py -m gpu -x -q > gpurun_out/s36_parity_tests.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/s36_parity_tests.log
timeout 600 python bench.py --no-cpu-baseline --no-extras > gpurun_out/s36_bench.json 2> gpurun_out/s36_bench.err; echo "bench rc=$?"
