#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_chain_gpu.py -x -q 2>&1 | tail -3
timeout 900 python bench.py > gpurun_out/k_bench.json 2> gpurun_out/k_bench.err; echo "bench rc=$?"; tail -c 300 gpurun_out/k_bench.err
python - <<'PY'
import json
d=json.load(open('gpurun_out/k_bench.json'))
print(d['value'], d['ms_per_step'])
print(json.dumps(d['gptq_act_order']['decode_step']))
PY
