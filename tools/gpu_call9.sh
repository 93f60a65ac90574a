#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_chain_gpu.py -q -x 2>&1 | tail -2
timeout 900 python bench.py > gpurun_out/k_bench.json 2> gpurun_out/k_bench.err; echo "bench rc=$?"; tail -c 300 gpurun_out/k_bench.err
python - <<'PY'
import json
d=json.load(open('gpurun_out/k_bench.json'))
print(d['value'], d['ms_per_step'], d['per_linear_api']['ms_per_step'], d['e2e']['value'])
print(json.dumps(d.get('chain_m')))
print([ (p.get('M'), p.get('ms_per_step')) for p in d['layer_chain']['points']])
PY
