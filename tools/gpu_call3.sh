#!/bin/bash
mkdir -p gpurun_out
rm -f gpurun_out/c9_*.log
for c in 1 0; do
  echo "== NO_CLUSTER=$c" >> gpurun_out/c9_pf.log
  W4A16_PF_NO_CLUSTER=$c timeout 300 python tools/prefill_check.py --shapes qkv o gate_up down --m 512 4096 4100 8192 --time >> gpurun_out/c9_pf.log 2>&1
done
cat gpurun_out/c9_pf.log
timeout 900 python -m pytest tests/test_config3_prefill_gpu.py -m gpu -x -q > gpurun_out/c9_pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/c9_pytest.log
tail -5 gpurun_out/c9_pytest.log
