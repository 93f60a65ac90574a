#!/bin/bash
timeout 600 python -m pytest tests/test_tp_chain_gpu.py -x -q 2>&1 | tail -30
