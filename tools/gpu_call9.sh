#!/bin/bash
mkdir -p gpurun_out
timeout 1800 python -m pytest tests -m gpu -q > gpurun_out/l_all.log 2>&1
echo "pytest rc=$?" >> gpurun_out/l_all.log
tail -12 gpurun_out/l_all.log
