#!/bin/bash
mkdir -p gpurun_out
timeout 120 python tools/prof_chain.py > gpurun_out/j_prof_plain.log 2>&1 && \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:chain_kernel -s 1 -c 1 -o gpurun_out/j_chain2 -f python tools/prof_chain.py > gpurun_out/j_ncu.log 2>&1
tail -3 gpurun_out/j_ncu.log
