#!/bin/bash
mkdir -p gpurun_out
python scripts/gpu_dbg_scaling.py scaling_clades_a002_r4 2>&1 | tail -30
python -m pytest tests/test_gpu_newick.py -x -q -m gpu 2>&1 | tail -3
python scripts/gpu_newick_probe.py c4 16384 2>&1 | grep '"kernel": 2' | tail -1
python -m pytest tests/test_scaling_regime.py -q -m gpu 2>&1 | tail -25
