#!/bin/bash
# launch list of the bench command (metrics-only ncu pass, after the same command has exited 0 without ncu)
mkdir -p gpurun_out
timeout 400 python bench.py --quick --steps 2 --warmup 3 > gpurun_out/q_bench.json 2> gpurun_out/q_bench.err || { echo "plain run failed"; tail -3 gpurun_out/q_bench.err; exit 1; }
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"chain_kernel|gemv_imma|decode_kernel|prefill_kernel|gemv_stream|midm_kernel" -c 900 --csv --log-file gpurun_out/r2_bench_launches.csv python bench.py --quick --steps 2 --warmup 3 > gpurun_out/q_ncu.log 2>&1
echo "ncu rc=$?"; wc -l gpurun_out/r2_bench_launches.csv
