#!/bin/bash
mkdir -p gpurun_out
N=$1
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29551 bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/v_tp${N}.json 2> gpurun_out/v_tp${N}.err
echo "rc=$?"; grep -c '^{' gpurun_out/v_tp${N}.json; grep -i "error\|Traceback" gpurun_out/v_tp${N}.err | head -5
