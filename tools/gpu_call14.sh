#!/bin/bash
mkdir -p gpurun_out
N=$1
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29541 bench.py --gpus $N --steps 20 --warmup 5 --quick > gpurun_out/r_tp${N}_chain.json 2> gpurun_out/r_tp${N}_chain.err
echo "rc=$?"; grep -o "\"n_gpus\": $N, \"steps\": 20, \"warmup\": 5, \"ms_per_step\": [0-9.]*" gpurun_out/r_tp${N}_chain.json; grep -o '"chain": {[^}]*}' gpurun_out/r_tp${N}_chain.json; grep -o '"launch": "[^"]*"' gpurun_out/r_tp${N}_chain.json; grep -i "error\|unavailable" gpurun_out/r_tp${N}_chain.err | head -5
