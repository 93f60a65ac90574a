#!/bin/bash
mkdir -p gpurun_out
# 1. launch list of the bench command itself (after it exited 0 without ncu)
timeout 600 python bench.py --quick --steps 2 --warmup 3 > gpurun_out/t_plain.json 2> gpurun_out/t_plain.err && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"chain_kernel|gemv_imma|decode_kernel|prefill_kernel|gemv_stream" -c 900 --csv --log-file gpurun_out/r2_bench_launches.csv python bench.py --quick --steps 2 --warmup 3 > gpurun_out/t_ncu.log 2>&1
echo "launch list rc=$?"; wc -l gpurun_out/r2_bench_launches.csv
# 2. prefill at M = 4096, full set
timeout 300 python tools/prof_target.py --m 4096 --copies 1 --reps 2 > gpurun_out/t_pf_plain.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:prefill_kernel -s 4 -c 4 -o gpurun_out/r2_prefill -f python tools/prof_target.py --m 4096 --copies 1 --reps 2 > gpurun_out/t_pf_ncu.log 2>&1
echo "prefill ncu rc=$?"; tail -2 gpurun_out/t_pf_ncu.log
