#!/bin/bash
# GPU session 3 of round 2: ordered split-K of the GEMM tail wave (correctness, determinism, timing), the new parity
# tests against the reference's whole forward, the bench line with its new keys, a launch list of one DPT head.
cd "$(dirname "$0")/.." || exit 1
mkdir -p gpurun_out
ST=tests/native/selftest
L=gpurun_out/s3_selftest.log
: > $L
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw,temperature.gpu --format=csv >> $L 2>&1
run() { timeout 180 $ST "$@" >> $L 2>&1; rc=$?; [ $rc -ne 0 ] && echo "  -> rc=$rc for: $*" >> $L; return 0; }
echo "== GEMM tail split (ordered reduce-add chain)" >> $L
run gemm_resid 11008 1024 1024 1 0 37 1
run gemm_resid 11008 1024 4096 1 0 37 1
run gemm_resid 300 1024 192 1 0 1 1
run gemm_resid 5000 1024 1024 0 0 13 1
run gemm_resid 22016 1024 4096 1 0 97 1
run gemm_resid 1370 1024 1024 1 0 7 1
run gemm_resid 11008 1024 1024 1 1 37 1
for shape in "11008 1024 1024" "11008 1024 4096" "5504 1024 4096" "22016 1024 1024" "22016 1024 4096" "10992 1024 1024" "10992 1024 4096"; do
  run bench_gemm 3 $shape
  run bench_gemm 4 $shape
done
grep -E "BENCH|FAIL|rc=|PASS" $L | tail -60

echo "== GPU tests"
timeout 1800 python -m pytest tests -m gpu -x -q > gpurun_out/s3_gpu_tests.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/s3_gpu_tests.log
echo "== bench (all keys)"
timeout 1200 python bench.py --steps 10 --warmup 3 > gpurun_out/s3_bench_default.json 2> gpurun_out/s3_bench_default.err; echo "bench rc=$?"; tail -3 gpurun_out/s3_bench_default.err
timeout 900 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-extras --set gemm_split=0 > gpurun_out/s3_bench_nosplit.json 2> gpurun_out/s3_bench_nosplit.err; echo "bench nosplit rc=$?"
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/s3_bench_*.json')):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1])
        print(f, round(d['value'],1), 'e2e', round(d['e2e']['value'],1), 'attn TF', round(d['roofline']['achieved'],1), {k:v['ms'] for k,v in d['breakdown_ms'].items()})
        for k in ('roofline_general','strong','torch_gpu_baseline','cpu_baseline'):
            if k in d: print('  ', k, json.dumps(d[k])[:600])
        print('   kernels', json.dumps(d['roofline'].get('kernels'))[:1500])
    except Exception as e:
        print(f, 'unreadable', e)
PY
echo "== launch list of one DPT head + camera head"
timeout 300 python tools/profile_heads.py > gpurun_out/s3_profile_heads.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/s3_heads_launches.csv python tools/profile_heads.py > gpurun_out/s3_ncu_heads.log 2>&1
echo "ncu heads rc=$?"; cat gpurun_out/s3_profile_heads.log | tail -3
