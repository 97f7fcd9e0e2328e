#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests/test_gpu_newick.py -x -q -m gpu 2>&1 | tail -1
for args in "" "--batch 2048 --opt host_threads=4"; do
  python bench.py --no-extras --no-cpu-baseline $args > gpurun_out/bench_c4_r4k.json 2> gpurun_out/bench_c4_r4k.err; tail -2 gpurun_out/bench_c4_r4k.err
  python -c "
import json; d=json.loads(open('gpurun_out/bench_c4_r4k.json').read().splitlines()[-1]); print('[$args]', d['value'], 'e2e', d['e2e']['value'], json.dumps(d['e2e']['last_call_ms_rank0'])); print('   arrays', d['e2e_arrays']['value'], d['e2e_arrays']['last_call_ms_rank0']['ms_call'])"
done
