#!/bin/bash
mkdir -p gpurun_out
{
timeout 200 python tools/midm_sweep.py --check --dispatch auto prefill --m 40 48 64 --shapes qkv o gate_up down 2>&1 | tail -24
} > gpurun_out/midm_64.log 2>&1
cat gpurun_out/midm_64.log
