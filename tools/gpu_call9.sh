#!/bin/bash
mkdir -p gpurun_out
timeout 1800 python -m pytest tests -m gpu -q -x > gpurun_out/l_all.log 2>&1
echo "pytest rc=$?" >> gpurun_out/l_all.log
tail -4 gpurun_out/l_all.log
timeout 900 python bench.py --quick > gpurun_out/s_bench.json 2> gpurun_out/s_bench.err; echo "bench rc=$?"
grep -o '"value": [0-9.]*, "unit": "GB/s", "n_gpus": 1, "steps": [0-9]*, "warmup": [0-9]*, "ms_per_step": [0-9.]*' gpurun_out/s_bench.json
