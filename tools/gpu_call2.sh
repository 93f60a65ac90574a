#!/bin/bash
mkdir -p gpurun_out
timeout 300 ./tools/ubench/l2pf > gpurun_out/c2_l2pf.log 2>&1
for sh in gate_up down; do
  W4A16_L2PF=5 timeout 120 python tools/gemv_timeline.py --shape $sh --graph --next > gpurun_out/c2_tl_${sh}_pf5.log 2>&1
  W4A16_L2PF=1 timeout 120 python tools/gemv_timeline.py --shape $sh --graph --next > gpurun_out/c2_tl_${sh}_pf1.log 2>&1
done
cat gpurun_out/c2_l2pf.log
