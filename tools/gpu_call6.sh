#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_chain_gpu.py -x -q > gpurun_out/h_chain_test.log 2>&1
tail -25 gpurun_out/h_chain_test.log
timeout 300 python tools/decode_ab.py --api chain --chain --cfg "" 2>&1 | grep "\[ab\]" | tee gpurun_out/h_ab.log
timeout 300 python tools/decode_ab.py --api chain --cfg "" 2>&1 | grep "\[ab\]" | tee -a gpurun_out/h_ab.log
timeout 300 python tools/chain_timeline.py > gpurun_out/h_ctl.log 2>&1
cat gpurun_out/h_ctl.log
