#!/bin/bash
python -m pytest tests/test_gpu_phylohmm.py -x -q -m gpu 2>&1 | tail -2
LH_TRACE=1 python scripts/gpu_pipe_probe.py 16384 2>&1 | grep "batch 256\|RunPipeline busy" | sed -n 3,5p
python scripts/gpu_pipe_probe.py 4096 2>&1 | grep "batch 256" | tail -2
