#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 20 --warmup 5 > gpurun_out/m_tp2.json 2> gpurun_out/m_tp2.err
echo "tp2 rc=$?"; tail -c 400 gpurun_out/m_tp2.err; head -c 1200 gpurun_out/m_tp2.json
timeout 600 python -m pytest tests/test_tp_fused_gpu.py -x -q 2>&1 | tail -3
