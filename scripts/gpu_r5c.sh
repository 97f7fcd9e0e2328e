#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests/test_gpu_newick.py -x -q -m gpu -k "3 or lenient or chunks" 2>&1 | tail -3
python scripts/gpu_newick_probe.py c4 2048 2>&1 | grep -v '"kernel": 1' | tail -6 | cut -c1-120
scripts/gpu_ncu_newick.sh k_newick_warp r5c | cut -c1-120
