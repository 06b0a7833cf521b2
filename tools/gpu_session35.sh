#!/bin/bash
# Residual prefetch in the convolution epilogue: head parity tests, per-launch times of one head, the bench line.
cd "$(dirname "$0")/.." || exit 1
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_heads_gpu.py tests/test_gs_gpu.py -m gpu -q -x > gpurun_out/s35_head_tests.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/s35_head_tests.log
timeout 300 python tools/profile_heads.py > gpurun_out/s35_profile_heads.log 2>&1; tail -2 gpurun_out/s35_profile_heads.log
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/s35_heads_launches.csv python tools/profile_heads.py > gpurun_out/s35_ncu_heads.log 2>&1; echo "ncu heads rc=$?"
timeout 600 python bench.py --no-cpu-baseline --no-extras > gpurun_out/s35_bench.json 2> gpurun_out/s35_bench.err; echo "bench rc=$?"
