#!/bin/bash
# last check of the in-tree build: smoke + the GPU suite
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -2
