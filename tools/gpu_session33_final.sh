#!/bin/bash
# Final single-GPU validation of round 2: whole GPU suite, smoke, the bench line with all keys, then (after both exited 0
# without a profiler) the ncu launch list of the bench command and one --set full capture of the attention kernel that
# chooses its softmax form on the device, at the DINOv2 shape.
cd "$(dirname "$0")/.." || exit 1
mkdir -p gpurun_out
timeout 1800 python -m pytest tests -m gpu -q > gpurun_out/s33_gpu_tests.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/s33_gpu_tests.log
timeout 600 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
timeout 1500 python bench.py > gpurun_out/s33_bench.json 2> gpurun_out/s33_bench.err; echo "bench rc=$?"; tail -2 gpurun_out/s33_bench.err
timeout 1200 ncu --metrics gpu__time_duration.sum --clock-control none -s 2500 -c 2000 --csv --log-file gpurun_out/r02_final_launches.csv \
  python bench.py --steps 1 --warmup 0 --no-cpu-baseline --no-extras > gpurun_out/s33_ncu_bench.log 2>&1; echo "ncu launch list rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:attn_split_kernel -s 6 -c 1 -f -o gpurun_out/r02_attn_dyn_dino \
  python tools/bench_guard.py 8 1374 "one launch, data fits" > gpurun_out/s33_ncu_attn.log 2>&1; echo "ncu full rc=$?"
ls -la gpurun_out/r02_attn_dyn_dino.ncu-rep gpurun_out/r02_final_launches.csv
