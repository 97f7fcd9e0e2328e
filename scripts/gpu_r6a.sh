#!/bin/bash
for rs in "8 8" "12 4" "14 4" "16 4" "12 8" "16 8" "10 6"; do
  set -- $rs
  echo -n "reader $1 sampler $2: "; LH_READER_THREADS=$1 LH_SAMPLER_THREADS=$2 python scripts/gpu_pipe_probe.py 16384 2>&1 | grep "batch 256" | tail -1
done
echo -n "default split, 4096 rows: "; python scripts/gpu_pipe_probe.py 4096 2>&1 | grep "batch 256" | tail -1
