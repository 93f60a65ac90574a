#!/bin/bash
mkdir -p gpurun_out
{
timeout 200 python tools/midm_sweep.py --check --dispatch auto decode --m 9 10 12 14 --shapes qkv o gate_up down 2>&1 | tail -32
} > gpurun_out/midm_low.log 2>&1
cat gpurun_out/midm_low.log
