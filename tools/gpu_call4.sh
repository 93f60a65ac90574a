#!/bin/bash
mkdir -p gpurun_out
rm -f gpurun_out/d2_*.log
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/d2_all.log 2>&1
echo "pytest rc=$?" >> gpurun_out/d2_all.log
tail -30 gpurun_out/d2_all.log
