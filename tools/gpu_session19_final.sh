#!/bin/bash
# Final single-GPU validation of the round: whole GPU test suite, smoke, the bench line with all keys.
cd "$(dirname "$0")/.." || exit 1
mkdir -p gpurun_out
timeout 1800 python -m pytest tests -m gpu -q > gpurun_out/s19_gpu_tests.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/s19_gpu_tests.log
timeout 600 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
timeout 1500 python bench.py --steps 20 --warmup 5 > gpurun_out/s19_bench.json 2> gpurun_out/s19_bench.err; echo "bench rc=$?"; tail -3 gpurun_out/s19_bench.err
python - <<'PY'
import json
d=json.loads(open('gpurun_out/s19_bench.json').read().strip().splitlines()[-1])
print(round(d['value'],1), 'ms', round(d['ms_per_step'],2), 'e2e', round(d['e2e']['value'],1), 'attn TF', round(d['roofline']['achieved'],1), 'frac', round(d['roofline']['frac'],3), d['gpu_launches'], d['clocks'])
print({k:v['ms'] for k,v in d['breakdown_ms'].items()})
for k in ('roofline_general','strong','cpu_baseline'):
    print(k, json.dumps(d.get(k))[:400])
print('torch', json.dumps(d.get('torch_gpu_baseline'))[:700])
print([ (x['class'], x['achieved'], x['frac']) for x in d['roofline']['kernels']])
PY
