#!/bin/bash
timeout 900 python -m pytest tests/test_chain_gpu.py -x -q 2>&1 | tail -2
timeout 300 python tools/decode_ab.py --api chain --chain --cfg "" --cfg "A=1" 2>&1 | grep "\[ab\] "
timeout 300 python tools/chain_timeline.py 2>&1 | grep -A8 "^-- o"
