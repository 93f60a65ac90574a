#!/bin/bash
mkdir -p gpurun_out
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -3
timeout 1800 python -m pytest tests -m gpu -q > gpurun_out/l_all.log 2>&1
echo "pytest rc=$?" >> gpurun_out/l_all.log
tail -3 gpurun_out/l_all.log
timeout 900 python bench.py > gpurun_out/k_bench.json 2> gpurun_out/k_bench.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.load(open('gpurun_out/k_bench.json'))
print(d['value'], d['ms_per_step'], d['roofline']['frac'], d['per_linear_api']['ms_per_step'], d['e2e']['value'], d['prefill']['M4096_TFLOPS'])
PY
