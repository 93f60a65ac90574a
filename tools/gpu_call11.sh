#!/bin/bash
mkdir -p gpurun_out
run() { # name, env...
  name=$1; shift
  env "$@" timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29521 bench.py --gpus 8 --steps 20 --warmup 5 --quick > gpurun_out/p_tp8_$name.json 2> gpurun_out/p_tp8_$name.err
  echo "$name rc=$?"; grep -o '"value": [0-9.]*, "unit": "GB/s", "n_gpus": 8, "steps": 20, "warmup": 5, "ms_per_step": [0-9.]*' gpurun_out/p_tp8_$name.json
}
run base A=0
run pdlattr W4A16_FUSED_PDL_ATTR=1
run early W4A16_FUSED_PDL_ATTR=1 W4A16_FUSED_EARLY_PDL=1
