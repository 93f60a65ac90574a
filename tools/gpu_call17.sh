#!/bin/bash
mkdir -p gpurun_out
{
timeout 900 python -m pytest tests/test_config3_prefill_gpu.py tests/test_parity_gpu.py tests/test_fused_gpu.py -x -q 2>&1 | tail -2
timeout 300 python tools/midm_sweep.py --check --dispatch auto --m 2048 4096 64 128 256 --shapes gate_up qkv o down 2>&1 | tail -20
} > gpurun_out/pf_cl2pdl.log 2>&1
cat gpurun_out/pf_cl2pdl.log
