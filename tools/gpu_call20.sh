#!/bin/bash
# 2-GPU check: the multi-GPU tests and a short tensor-parallel bench line
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -k "tp or peer or rank or fused_ar or allreduce or row" > gpurun_out/tp2_tests.log 2>&1; echo "pytest rc=$?" >> gpurun_out/tp2_tests.log
tail -3 gpurun_out/tp2_tests.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 5 --warmup 3 > gpurun_out/tp2_bench.json 2> gpurun_out/tp2_bench.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.loads(open('gpurun_out/tp2_bench.json').read().strip().splitlines()[-1])
print(d['metric'], d['value'], d['unit'], d['ms_per_step'], d.get('e2e'))
PY
