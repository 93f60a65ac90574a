#!/bin/bash
# the static-scratch entry point: alone in a fresh process, then after the other GPU tests
mkdir -p gpurun_out
{
echo "== alone"; timeout 600 python -m pytest tests/test_zz_no_alloc.py -x -q 2>&1 | tail -3
echo "== whole suite"; timeout 2400 python -m pytest tests -x -q -m gpu 2>&1 | tail -3
} > gpurun_out/no_alloc.log 2>&1
cat gpurun_out/no_alloc.log
