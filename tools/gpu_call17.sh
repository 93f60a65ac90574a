#!/bin/bash
# ship build: full GPU tests, then the mid-M sweep (new route against forced prefill / decode on the same box)
mkdir -p gpurun_out
{
timeout 1500 python -m pytest tests -x -q -m gpu 2>&1 | tail -3
echo "== auto"; timeout 200 python tools/midm_sweep.py --check --dispatch auto --m 12 13 14 16 17 24 32 33 --shapes qkv o gate_up down 2>&1 | tail -32
echo "== decode"; timeout 200 python tools/midm_sweep.py --dispatch decode --m 14 16 --shapes qkv o gate_up down 2>&1 | tail -8
echo "== prefill"; timeout 200 python tools/midm_sweep.py --dispatch prefill --m 17 24 32 --shapes qkv o gate_up down 2>&1 | tail -12
} > gpurun_out/midm_final.log 2>&1
cat gpurun_out/midm_final.log
