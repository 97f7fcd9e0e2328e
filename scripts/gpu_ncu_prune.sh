#!/bin/bash
# One full ncu capture of k_prune (after the same command ran clean without ncu). usage: scripts/gpu_ncu_prune.sh <tag>
tag=${1:-rX}
mkdir -p gpurun_out
python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-e2e --batch 4096 > gpurun_out/plain2_$tag.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:k_prune -s 3 -c 1 -o gpurun_out/prof_prune_$tag -f python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-e2e --batch 4096 > gpurun_out/ncu_full_$tag.log 2>&1
tail -3 gpurun_out/ncu_full_$tag.log
