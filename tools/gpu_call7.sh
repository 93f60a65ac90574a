#!/bin/bash
timeout 400 python tools/decode_ab.py --cfg "" --cfg "W4A16_GV_L2PF=2" --cfg "W4A16_GV_L2PF=4" --cfg "W4A16_GV_L2PF=8" --cfg "W4A16_GV_L2PF=16" 2>&1 | grep "\[ab\] "
