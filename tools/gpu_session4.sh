#!/bin/bash
# GPU session 4 of round 2: K-split chain length experiments, DPT tail chunking, new tests.
cd "$(dirname "$0")/.." || exit 1
mkdir -p gpurun_out
ST=tests/native/selftest
L=gpurun_out/s4_selftest.log
: > $L
run() { timeout 180 "$@" >> $L 2>&1; rc=$?; [ $rc -ne 0 ] && echo "  -> rc=$rc for: $*" >> $L; return 0; }
for ms in 2 3 4; do
  echo "== WM_GEMM_MAX_SPLIT=$ms WM_GEMM_SPLIT_MIN_KB=1" >> $L
  for shape in "11008 1024 1024" "11008 1024 4096" "5504 1024 4096" "5504 1024 1024"; do
    run env WM_GEMM_MAX_SPLIT=$ms WM_GEMM_SPLIT_MIN_KB=1 $ST bench_gemm 4 $shape
  done
done
echo "== defaults" >> $L
for shape in "11008 1024 1024" "11008 1024 4096"; do run $ST bench_gemm 3 $shape; run $ST bench_gemm 4 $shape; done
grep -E "==|BENCH|FAIL|rc=" $L
echo "== DPT tail views"
for tv in 8 4 2 1; do
  echo "tail_views=$tv"; WM_DPT_TAIL_VIEWS=$tv timeout 300 python tools/profile_heads.py 2>&1 | tail -1
done
echo "== tests"
timeout 1500 python -m pytest tests/test_heads_gpu.py tests/test_gs_gpu.py tests/test_parity_gpu.py -m gpu -x -q -k "not full_length" > gpurun_out/s4_gpu_tests.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/s4_gpu_tests.log
echo "== bench"
for tv in 8 2; do
WM_DPT_TAIL_VIEWS=$tv timeout 900 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-extras > gpurun_out/s4_bench_tv$tv.json 2> gpurun_out/s4_bench_tv$tv.err; echo "bench tv=$tv rc=$?"
done
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/s4_bench_*.json')):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1])
        print(f, round(d['value'],1), 'e2e', round(d['e2e']['value'],1), 'attn TF', round(d['roofline']['achieved'],1), {k:v['ms'] for k,v in d['breakdown_ms'].items()})
    except Exception as e:
        print(f, 'unreadable', e)
PY
echo "== ncu of the split proj GEMM"
NCU="ncu --set full --clock-control none --import-source on"
WM_GEMM_SPLIT_MIN_KB=1 timeout 300 $ST bench_gemm 4 11008 1024 1024 > gpurun_out/ncu_plain_proj.log 2>&1 && \
WM_GEMM_SPLIT_MIN_KB=1 timeout 600 $NCU -k regex:gemm_kernel -s 5 -c 1 -f -o gpurun_out/s4_gemm_proj_chain $ST bench_gemm 4 11008 1024 1024 > gpurun_out/ncu_proj.log 2>&1
echo "ncu rc=$?"
