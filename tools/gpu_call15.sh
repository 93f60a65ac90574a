#!/bin/bash
mkdir -p gpurun_out
timeout 300 python tools/chain_timeline.py > gpurun_out/u_ctl.log 2>&1; tail -45 gpurun_out/u_ctl.log
timeout 120 python tools/prof_chain.py > gpurun_out/u_prof_plain.log 2>&1 && \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:chain_kernel -s 1 -c 1 -o gpurun_out/u_chain -f python tools/prof_chain.py > gpurun_out/u_ncu.log 2>&1
tail -2 gpurun_out/u_ncu.log
