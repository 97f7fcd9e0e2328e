#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests/test_gpu_newick.py tests/test_zz_no_alloc.py -x -q -m gpu 2>&1 | tail -3
python scripts/gpu_newick_probe.py c4 148 2>&1 | grep -v '"kernel": 1' | tail -6 | cut -c1-110
python scripts/gpu_newick_probe.py c4 4096 2>&1 | grep -v '"kernel": 1' | tail -6 | cut -c1-110
scripts/gpu_r5f.sh 2>&1 | head -3
grep "dev compute chunk" gpurun_out/trace_r5f.err | tail -10
python bench.py --no-extras --no-cpu-baseline > gpurun_out/bench_c4_r5g.json 2> gpurun_out/bench_c4_r5g.err; tail -2 gpurun_out/bench_c4_r5g.err
python -c "
import json; d=json.loads(open('gpurun_out/bench_c4_r5g.json').read().splitlines()[-1]); print(d['value'], 'e2e', d['e2e']['value'], json.dumps(d['e2e']['last_call_ms_rank0'])); print('   arrays', d['e2e_arrays']['value'], d['e2e_arrays']['last_call_ms_rank0']['ms_call'])"
