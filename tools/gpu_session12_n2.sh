#!/bin/bash
# 2-GPU session: partitioned voxel merge (bit-equality against the single-GPU merge), bench at N = 2 with and without it.
cd "$(dirname "$0")/.." || exit 1
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 400 $TR --master-port 29561 tests/sp_check.py > gpurun_out/s12_sp_check.log 2>&1; echo "sp_check rc=$?"; grep -E "rank 0|SP_CHECK|Error|error|Traceback" gpurun_out/s12_sp_check.log | tail -24
timeout 500 $TR --master-port 29562 bench.py --gpus 2 --steps 10 --warmup 3 --no-extras > gpurun_out/s12_bench_n2.json 2> gpurun_out/s12_bench_n2.err; echo "bench rc=$?"; tail -3 gpurun_out/s12_bench_n2.err
WM_SPLAT_PARTITION=0 timeout 500 $TR --master-port 29563 bench.py --gpus 2 --steps 10 --warmup 3 --no-extras --no-sp-parity > gpurun_out/s12_bench_n2_replicated.json 2> gpurun_out/s12_bench_n2_replicated.err; echo "bench (replicated merge) rc=$?"
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/s12_bench_*.json')):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1])
        print(f, round(d['value'],1), 'ms', round(d['ms_per_step'],2), 'e2e', round(d['e2e']['value'],1), d['e2e']['d2h_bytes_per_step'], 'attn TF', round(d['roofline']['achieved'],1), {k:v['ms'] for k,v in d['breakdown_ms'].items() if k in ('camera_head','splats','dpt_gaussian')})
    except Exception as e:
        print(f, 'unreadable', e)
PY
