#!/bin/bash
# final validation of the round: smoke, GPU tests, the default bench line, a mid-M sweep, and an ncu capture of the mid-M kernel
mkdir -p gpurun_out
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -3
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/l_all.log 2>&1
echo "pytest rc=$?" >> gpurun_out/l_all.log
tail -3 gpurun_out/l_all.log
timeout 900 python bench.py > gpurun_out/k_bench.json 2> gpurun_out/k_bench.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.load(open('gpurun_out/k_bench.json'))
print(d['value'], d['ms_per_step'], d['roofline']['frac'], d['per_linear_api']['ms_per_step'], d['e2e']['value'], d['prefill']['M4096_TFLOPS'])
PY
timeout 200 python tools/midm_sweep.py --check --dispatch auto --m 15 16 24 32 --shapes qkv o gate_up down > gpurun_out/midm_auto.log 2>&1; tail -16 gpurun_out/midm_auto.log
timeout 200 python tools/midm_sweep.py --dispatch prefill --m 24 32 --shapes qkv o gate_up down > gpurun_out/midm_prefill.log 2>&1; tail -8 gpurun_out/midm_prefill.log
timeout 200 python tools/midm_sweep.py --dispatch decode --m 15 16 --shapes qkv o gate_up down > gpurun_out/midm_decode.log 2>&1; tail -8 gpurun_out/midm_decode.log
timeout 120 python tools/prof_target.py --m 32 --shapes gate_up --copies 3 --reps 2 > gpurun_out/pt32.log 2>&1 && \
timeout 300 ncu --set full --clock-control none --import-source on -k regex:midm_kernel -s 3 -c 1 -o gpurun_out/midm_m32 -f python tools/prof_target.py --m 32 --shapes gate_up --copies 3 --reps 2 > gpurun_out/ncu_midm32.log 2>&1
tail -1 gpurun_out/ncu_midm32.log
