#!/bin/bash
# Final 2-GPU validation: the pytest entry that needs two GPUs, then the bench line at N = 2.
cd "$(dirname "$0")/.." || exit 1
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_sp_gpu.py -m gpu -q -s > gpurun_out/s32_sp_tests.log 2>&1; echo "pytest rc=$?"; grep -E "rank 0|SP_CHECK|passed|failed" gpurun_out/s32_sp_tests.log | tail -20
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 900 $TR --master-port 29583 bench.py --gpus 2 --steps 20 --warmup 5 > gpurun_out/s32_bench_n2.json 2> gpurun_out/s32_bench_n2.err; echo "bench rc=$?"; tail -2 gpurun_out/s32_bench_n2.err
python - <<'PY'
import json
d=json.loads(open('gpurun_out/s32_bench_n2.json').read().strip().splitlines()[-1])
print(round(d['value'],1), 'ms', round(d['ms_per_step'],2), 'e2e', round(d['e2e']['value'],1), 'attn TF', round(d['roofline']['achieved'],1), {k:v['ms'] for k,v in d['breakdown_ms'].items()})
for k in ('sp_parity','strong'): print(k, json.dumps(d.get(k))[:300])
PY
