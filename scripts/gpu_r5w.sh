#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests/test_gpu_phylohmm.py tests/test_gpu_catch_compat.py -x -q -m gpu 2>&1 | tail -1
python bench.py --config c3 --no-cpu-baseline > gpurun_out/bench_c3_r5w.json 2> gpurun_out/bench_c3_r5w.err; tail -1 gpurun_out/bench_c3_r5w.err | cut -c1-200
python -c "
import json; d=json.loads(open('gpurun_out/bench_c3_r5w.json').read().splitlines()[-1]); print(d['value'], 'e2e', d['e2e']['value'], 'arrays', d['e2e_arrays']['value']); print(d['e2e_pipeline'])"
python bench.py --no-extras --no-cpu-baseline > gpurun_out/bench_c4_r5w.json 2> gpurun_out/bench_c4_r5w.err; tail -1 gpurun_out/bench_c4_r5w.err | cut -c1-200
python -c "
import json; d=json.loads(open('gpurun_out/bench_c4_r5w.json').read().splitlines()[-1]); print(d['value'], 'e2e', d['e2e']['value'], json.dumps(d['e2e']['last_call_ms_rank0'])); print('   arrays', d['e2e_arrays']['value'])"
