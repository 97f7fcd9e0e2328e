#!/bin/bash
LH_TRACE=1 python scripts/gpu_pipe_probe.py 16384 2>&1 | grep "batch 256\|RunPipeline busy" | head -6
LH_HOST_SAMPLER=1 LH_TRACE=1 python scripts/gpu_pipe_probe.py 16384 2>&1 | grep "batch 256\|RunPipeline busy" | head -6
