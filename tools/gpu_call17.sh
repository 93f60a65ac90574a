#!/bin/bash
mkdir -p gpurun_out
timeout 150 python tools/midm_sweep.py --dispatch auto prefill --m 40 48 --shapes qkv o gate_up down 2>&1 | tail -16 | awk '{print $1,$4,$5,$6,$7,$8}' > gpurun_out/midm_48.log
cat gpurun_out/midm_48.log
