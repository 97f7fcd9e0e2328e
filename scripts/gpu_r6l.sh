#!/bin/bash
python -m pytest tests/test_gpu_sample_paths.py tests/test_gpu_phylohmm.py tests/test_gpu_newick.py -x -q -m gpu 2>&1 | tail -2
python scripts/gpu_pipe_probe.py 16384 2>&1 | grep "batch 256" | tail -1
python scripts/gpu_pipe_probe.py 4096 2>&1 | grep "batch 256" | tail -1
