#!/bin/bash
python -m pytest tests/test_gpu_newick.py -x -q -m gpu -k "pinned" 2>&1 | tail -1
LH_TRACE=1 python scripts/gpu_pipe_probe.py 4096 2>&1 | grep -v "LHTRACE [a-z ]*[0-9.]* ms$" | grep "batch\|RunPipeline busy" | tail -24
