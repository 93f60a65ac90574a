#!/bin/bash
mkdir -p gpurun_out
timeout 600 python tools/decode_ab.py --cfg "W4A16_L2PF=0" --cfg "W4A16_L2PF=49" --cfg "W4A16_L2PF=49,W4A16_L2PF_HOPS=2" \
   --cfg "W4A16_L2PF=81" --cfg "W4A16_L2PF=81,W4A16_L2PF_HOPS=2" --cfg "W4A16_L2PF=0" > gpurun_out/c5_ab_m1.log 2>&1
grep "\[ab\]" gpurun_out/c5_ab_m1.log
