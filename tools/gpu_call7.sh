#!/bin/bash
timeout 500 python tools/decode_ab.py --api chain --chain --cfg "W4A16_CHAIN_L2PF=6" --cfg "W4A16_CHAIN_L2PF=8" --cfg "W4A16_CHAIN_L2PF=12" --cfg "W4A16_CHAIN_L2PF=16" --cfg "W4A16_CHAIN_L2PF=24" --cfg "W4A16_CHAIN_L2PF=48" 2>&1 | grep "\[ab\] W"
