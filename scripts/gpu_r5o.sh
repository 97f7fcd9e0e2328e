#!/bin/bash
# C5 under the host share of one rank of eight (4 cores): how many slots?
mkdir -p gpurun_out
for sl in 6 4 3 8; do
  taskset -c 0-3 python bench.py --config c5 --families 128 --slots $sl --no-cpu-baseline --steps 3 --warmup 1 > gpurun_out/c5_r5o_$sl.json 2> gpurun_out/c5_r5o_$sl.err
  python -c "
import json; d=json.loads(open('gpurun_out/c5_r5o_$sl.json').read().splitlines()[-1]); print('slots $sl (4 cores)', d['value'], d.get('ms_per_step'))"
done
python bench.py --config c5 --families 1024 --no-cpu-baseline --steps 3 --warmup 1 > gpurun_out/c5_r5o_full.json 2> gpurun_out/c5_r5o_full.err
python -c "
import json; d=json.loads(open('gpurun_out/c5_r5o_full.json').read().splitlines()[-1]); print('slots 6 (16 cores, 1024 families)', d['value'], d.get('ms_per_step'))"
