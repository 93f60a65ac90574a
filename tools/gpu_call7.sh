#!/bin/bash
for m in 2 4; do
timeout 300 python tools/decode_ab.py --m $m --api chain --chain --cfg "" 2>&1 | grep "\[ab\] "
timeout 300 python tools/decode_ab.py --m $m --cfg "" 2>&1 | grep "\[ab\] "
done
