#!/bin/bash
# 8-GPU session: the bench line at N = 8 (sp_parity, strong, 128-view key) on the final state.
cd "$(dirname "$0")/.." || exit 1
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1"
timeout 900 $TR --master-port 29551 bench.py --gpus 8 --steps 10 --warmup 3 > gpurun_out/s9_bench_n8.json 2> gpurun_out/s9_bench_n8.err; echo "bench rc=$?"; tail -4 gpurun_out/s9_bench_n8.err
python - <<'PY'
import json
try:
    d=json.loads(open('gpurun_out/s9_bench_n8.json').read().strip().splitlines()[-1])
    print(round(d['value'],1), 'ms', round(d['ms_per_step'],2), 'e2e', round(d['e2e']['value'],1), 'attn TF', round(d['roofline']['achieved'],1), {k:v['ms'] for k,v in d['breakdown_ms'].items()})
    for k in ('sp_parity','strong','roofline_general','c4_128_views'):
        if k in d: print('  ', k, json.dumps(d[k])[:500])
except Exception as e:
    print('unreadable', e)
PY
