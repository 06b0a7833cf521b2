#!/bin/bash
# GPU session 6: the whole GPU test suite on the final state, then the bench line.
cd "$(dirname "$0")/.." || exit 1
mkdir -p gpurun_out
timeout 1800 python -m pytest tests -m gpu -x -q > gpurun_out/s6_gpu_tests.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/s6_gpu_tests.log
timeout 600 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -3
timeout 1200 python bench.py --steps 20 --warmup 5 > gpurun_out/s6_bench.json 2> gpurun_out/s6_bench.err; echo "bench rc=$?"; tail -3 gpurun_out/s6_bench.err
python - <<'PY'
import json
d=json.loads(open('gpurun_out/s6_bench.json').read().strip().splitlines()[-1])
print(round(d['value'],1), 'e2e', round(d['e2e']['value'],1), 'attn TF', round(d['roofline']['achieved'],1), {k:v['ms'] for k,v in d['breakdown_ms'].items()})
PY
