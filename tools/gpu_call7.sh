#!/bin/bash
timeout 900 python -m pytest tests/test_chain_gpu.py -x -q 2>&1 | tail -12
timeout 300 python tools/decode_ab.py --api chain --chain --cfg "" 2>&1 | grep "\[ab\] "
