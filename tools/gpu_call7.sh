#!/bin/bash
mkdir -p gpurun_out
timeout 300 python tools/decode_ab.py --api chain --chain --cfg "" --cfg "W4A16_CHAIN_INFLIGHT=1" --cfg "W4A16_CHAIN_INFLIGHT=2" --cfg "W4A16_CHAIN_INFLIGHT=3" --cfg "W4A16_CHAIN_INFLIGHT=4" 2>&1 | grep "\[ab\]" | tee gpurun_out/i_ab.log
W4A16_CHAIN_INFLIGHT=2 timeout 300 python tools/chain_timeline.py > gpurun_out/i_ctl2.log 2>&1
cat gpurun_out/i_ctl2.log
