#!/bin/bash
# GPU call 1 of round 2: A/B of the L2 prefetch hint, timelines, prefill wait breakdown, then the whole GPU suite.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/c1_smi.log 2>&1
timeout 600 python tools/decode_ab.py --cfg "W4A16_L2PF=0" --cfg "W4A16_L2PF=1" --cfg "W4A16_L2PF=2" \
   --cfg "W4A16_L2PF=5" --cfg "W4A16_L2PF=6" --cfg "W4A16_L2PF=1,W4A16_L2PF_HOPS=2" --cfg "W4A16_L2PF=2,W4A16_L2PF_HOPS=2" \
   --cfg "W4A16_L2PF=0" > gpurun_out/c1_ab_m1.log 2>&1
timeout 300 python tools/decode_ab.py --m 4 --cfg "W4A16_L2PF=0" --cfg "W4A16_L2PF=1" --cfg "W4A16_L2PF=2" > gpurun_out/c1_ab_m4.log 2>&1
for sh in o qkv gate_up down; do
  W4A16_L2PF=0 timeout 120 python tools/gemv_timeline.py --shape $sh --graph > gpurun_out/c1_tl_${sh}_pf0.log 2>&1
done
W4A16_PF_DEBUG=4 timeout 300 python tools/prefill_check.py --shapes qkv o gate_up down --m 4096 --time > gpurun_out/c1_pfdbg.log 2>&1
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/c1_pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/c1_pytest.log
tail -5 gpurun_out/c1_pytest.log
cat gpurun_out/c1_ab_m1.log | grep "\[ab\]"
