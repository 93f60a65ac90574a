#!/bin/bash
mkdir -p gpurun_out
timeout 600 python tools/decode_ab.py --cfg "W4A16_NO_GEMV_IMMA2=1" --cfg "W4A16_NO_GEMV_IMMA2=0" --cfg "W4A16_IMMA2_MAXK=4096" --cfg "W4A16_NO_GEMV_IMMA2=1" --cfg "W4A16_NO_GEMV_IMMA2=0" --cfg "W4A16_IMMA2_MAXK=4096" > gpurun_out/c7_ab_m1.log 2>&1
grep "\[ab\]" gpurun_out/c7_ab_m*.log
timeout 120 python tools/gemv_timeline.py --shape down --graph > gpurun_out/c7_tl_down.log 2>&1
grep -A6 "launch +0" gpurun_out/c7_tl_down.log
