#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_chain_gpu.py tests/test_abi.py -x -q 2>&1 | tail -3
timeout 300 python tools/chain_timeline.py > gpurun_out/o_ctl.log 2>&1; cat gpurun_out/o_ctl.log
timeout 900 python bench.py > gpurun_out/k_bench.json 2> gpurun_out/k_bench.err; echo "bench rc=$?"; tail -c 300 gpurun_out/k_bench.err
timeout 200 python tools/prof_chain.py --layers 36 > gpurun_out/k_prof.log 2>&1 && GIT_HASH=23da471-dirty timeout 600 python tools/capture_chain_traffic.py > gpurun_out/r2_chain_traffic.json 2> gpurun_out/k_cap.err
cat gpurun_out/r2_chain_traffic.json; tail -3 gpurun_out/k_cap.err
