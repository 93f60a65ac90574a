#!/bin/bash
mkdir -p gpurun_out
{
for i in 1 2; do
echo "== pdl $i"; timeout 200 python tools/midm_sweep.py --dispatch auto --m 64 128 --shapes o down gate_up 2>&1 | tail -6
echo "== no pdl $i"; W4A16_NO_PDL=1 timeout 200 python tools/midm_sweep.py --dispatch auto --m 64 128 --shapes o down gate_up 2>&1 | tail -6
done
} > gpurun_out/pf_pdl2.log 2>&1
cat gpurun_out/pf_pdl2.log
