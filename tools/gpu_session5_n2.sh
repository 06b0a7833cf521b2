#!/bin/bash
# 2-GPU session: sequence-parallel checks (camera head on local views, sharded priors), bench at N = 2 with sp_parity.
cd "$(dirname "$0")/.." || exit 1
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
echo "== sp_check (eager sharded camera head)"
timeout 600 $TR --master-port 29541 tests/sp_check.py > gpurun_out/s5_sp_check.log 2>&1; echo "sp_check rc=$?"; grep -E "rank 0|SP_CHECK|Error|error" gpurun_out/s5_sp_check.log | tail -30
echo "== sp_check with the sharded camera head captured in a CUDA graph"
WM_TEST_CAM_GRAPH=1 timeout 300 $TR --master-port 29542 tests/sp_check.py > gpurun_out/s5_sp_check_graph.log 2>&1; echo "sp_check graph rc=$?"; grep -E "rank 0: sharded camera|SP_CHECK|Error|error" gpurun_out/s5_sp_check_graph.log | tail -10
echo "== bench N=2"
timeout 900 $TR --master-port 29543 bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/s5_bench_n2.json 2> gpurun_out/s5_bench_n2.err; echo "bench rc=$?"; tail -3 gpurun_out/s5_bench_n2.err
WM_CAMERA_SHARDED=0 timeout 900 $TR --master-port 29544 bench.py --gpus 2 --steps 10 --warmup 3 --no-extras --no-sp-parity > gpurun_out/s5_bench_n2_camrepl.json 2> gpurun_out/s5_bench_n2_camrepl.err; echo "bench (replicated camera head) rc=$?"
WM_CAM_GRAPH=2 timeout 900 $TR --master-port 29545 bench.py --gpus 2 --steps 10 --warmup 3 --no-extras --no-sp-parity > gpurun_out/s5_bench_n2_camgraph.json 2> gpurun_out/s5_bench_n2_camgraph.err; echo "bench (graph-captured sharded camera head) rc=$?"
timeout 900 $TR --master-port 29546 bench.py --gpus 2 --steps 5 --warmup 3 --priors --no-extras > gpurun_out/s5_bench_n2_priors.json 2> gpurun_out/s5_bench_n2_priors.err; echo "bench priors rc=$?"; tail -3 gpurun_out/s5_bench_n2_priors.err
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/s5_bench_*.json')):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1])
        print(f, round(d['value'],1), 'ms', round(d['ms_per_step'],2), 'e2e', round(d['e2e']['value'],1), 'attn TF', round(d['roofline']['achieved'],1), {k:v['ms'] for k,v in d['breakdown_ms'].items()})
        for k in ('sp_parity','strong','roofline_general'):
            if k in d: print('  ', k, json.dumps(d[k])[:500])
    except Exception as e:
        print(f, 'unreadable', e)
PY
