#!/bin/bash
mkdir -p gpurun_out
N=$1
run() { name=$1; shift
  env "$@" timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29531 bench.py --gpus $N --steps 20 --warmup 5 --quick $EXTRA > gpurun_out/q_tp${N}_$name.json 2> gpurun_out/q_tp${N}_$name.err
  echo "$name rc=$?"; grep -o "\"n_gpus\": $N, \"steps\": 20, \"warmup\": 5, \"ms_per_step\": [0-9.]*" gpurun_out/q_tp${N}_$name.json; grep -o '"chain": {[^}]*}' gpurun_out/q_tp${N}_$name.json; tail -c 300 gpurun_out/q_tp${N}_$name.err | grep -v "^$" | tail -3
}
run chain A=0
EXTRA="--api linear" run linear A=0
