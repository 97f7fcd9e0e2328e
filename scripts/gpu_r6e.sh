#!/bin/bash
# the 2048-tree shard of one rank of eight (4 host threads): does reserving a CTA slot for the parser help there?
mkdir -p gpurun_out
for pr in 0 1 2; do
  taskset -c 0-3 python bench.py --batch 2048 --opt host_threads=4 --opt parse_reserve=$pr --no-extras --no-cpu-baseline --steps 8 --warmup 3 > gpurun_out/r6e_$pr.json 2> gpurun_out/r6e_$pr.err
  python -c "
import json; d=json.loads(open('gpurun_out/r6e_$pr.json').read().splitlines()[-1]); print('reserve $pr: value', round(d['value']), 'e2e', round(d['e2e']['value']), 'ms_call', round(d['e2e']['last_call_ms_rank0']['ms_call'],2), 'arrays', round(d['e2e_arrays']['value']))"
done
