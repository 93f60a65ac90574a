#!/bin/bash
mkdir -p gpurun_out
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -6
timeout 1800 python -m pytest tests -m gpu -q -x > gpurun_out/l_all.log 2>&1
echo "pytest rc=$?" >> gpurun_out/l_all.log
tail -5 gpurun_out/l_all.log
