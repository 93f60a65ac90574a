#!/bin/bash
mkdir -p gpurun_out
{
for ns in 0 2 3 4 6 8 12 16; do echo "== ns=$ns"; W4A16_MM_NS=$ns timeout 100 python tools/midm_sweep.py --dispatch auto --m 16 32 --shapes qkv o gate_up down 2>&1 | tail -8 | awk '{print $1,$4,$7,$8}'; done
} > gpurun_out/midm_ns.log 2>&1
cat gpurun_out/midm_ns.log
