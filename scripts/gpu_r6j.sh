#!/bin/bash
for th in 1 2; do
LH_TRACE=1 taskset -c 0-3 python scripts/gpu_c5_trace.py 150 $th 2> gpurun_out/c5trace.err | tail -1
sed -n '/=== traced call ===/,$p' gpurun_out/c5trace.err | grep LHTRACE | head -40
done
