#!/bin/bash
cd "$(dirname "$0")/.." || exit 1
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1"
timeout 600 $TR --master-port 29571 bench.py --gpus 8 --steps 10 --warmup 3 --no-extras > gpurun_out/s13_bench_n8.json 2> gpurun_out/s13_bench_n8.err; echo "bench rc=$?"; tail -3 gpurun_out/s13_bench_n8.err
WM_SPLAT_PARTITION=0 timeout 600 $TR --master-port 29572 bench.py --gpus 8 --steps 10 --warmup 3 --no-extras --no-sp-parity > gpurun_out/s13_bench_n8_replicated.json 2> gpurun_out/s13_bench_n8_replicated.err; echo "bench (replicated) rc=$?"
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/s13_bench_*.json')):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1])
        print(f, round(d['value'],1), 'ms', round(d['ms_per_step'],2), 'e2e', round(d['e2e']['value'],1), d['e2e']['d2h_bytes_per_step'], 'attn TF', round(d['roofline']['achieved'],1), {k:v['ms'] for k,v in d['breakdown_ms'].items() if k in ('camera_head','splats','dpt_gaussian','attn_global')})
    except Exception as e:
        print(f, 'unreadable', e)
PY
