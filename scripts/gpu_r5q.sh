#!/bin/bash
# RunPipeline on c3 with the stage busy times (LH_TRACE), 4096 and 16384 rows
mkdir -p gpurun_out
LH_TRACE=1 python bench.py --config c3 --no-cpu-baseline --steps 2 --warmup 1 > gpurun_out/pipe_r5q.json 2> gpurun_out/pipe_r5q.err
grep "RunPipeline busy" gpurun_out/pipe_r5q.err | tail -4
python -c "
import json; d=json.loads(open('gpurun_out/pipe_r5q.json').read().splitlines()[-1]); print(d['e2e_pipeline'])"
nproc; grep -c processor /proc/cpuinfo
