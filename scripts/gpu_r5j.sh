#!/bin/bash
mkdir -p gpurun_out
for pr in 0 1 2; do
  python bench.py --no-extras --no-cpu-baseline --opt parse_reserve=$pr > gpurun_out/bench_c4_r5j_$pr.json 2> gpurun_out/bench_c4_r5j_$pr.err; tail -1 gpurun_out/bench_c4_r5j_$pr.err | cut -c1-200
  python -c "
import json; d=json.loads(open('gpurun_out/bench_c4_r5j_$pr.json').read().splitlines()[-1]); print('reserve $pr', d['value'], 'e2e', d['e2e']['value'], json.dumps(d['e2e']['last_call_ms_rank0'])); print('   arrays', d['e2e_arrays']['value'], d['e2e_arrays']['last_call_ms_rank0']['ms_call'])"
done
