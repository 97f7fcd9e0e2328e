#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests/test_gpu_newick.py tests/test_zz_no_alloc.py -x -q -m gpu 2>&1 | tail -2
python scripts/gpu_newick_probe.py c4 16384 2>&1 | grep '"kernel": 2' | tail -1
for r in 0 1 2; do
  python bench.py --no-extras --no-cpu-baseline --opt parse_reserve=$r > gpurun_out/bench_c4_r4i_$r.json 2> gpurun_out/bench_c4_r4i.err; tail -2 gpurun_out/bench_c4_r4i.err
  python -c "
import json; d=json.loads(open('gpurun_out/bench_c4_r4i_$r.json').read().splitlines()[-1]); print('reserve $r', d['value'], 'e2e', d['e2e']['value'], json.dumps(d['e2e']['last_call_ms_rank0'])); print('   arrays', d['e2e_arrays']['value'])"
done
