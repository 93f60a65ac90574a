#!/bin/bash
mkdir -p gpurun_out
{
timeout 600 python -m pytest tests/test_parity_gpu.py tests/test_fused_gpu.py -x -q 2>&1 | tail -2
timeout 200 python tools/midm_sweep.py --check --dispatch auto --m 17 24 25 32 33 40 41 48 --shapes qkv gate_up 2>&1 | tail -16
} > gpurun_out/midm_nt.log 2>&1
cat gpurun_out/midm_nt.log
