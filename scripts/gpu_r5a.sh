#!/bin/bash
# r5a: k_newick_warp - parity, standalone parser time (kernels 3/2/1 on C4 and C3), the C4 line with it
mkdir -p gpurun_out
python -m pytest tests/test_gpu_newick.py tests/test_zz_no_alloc.py -x -q -m gpu 2>&1 | tail -3
python scripts/gpu_newick_probe.py c4 2048 2>&1 | grep -v '"kernel": 1' | tail -6
python scripts/gpu_newick_probe.py c3 16384 2>&1 | tail -9
python bench.py --no-extras --no-cpu-baseline > gpurun_out/bench_c4_r5a.json 2> gpurun_out/bench_c4_r5a.err; tail -2 gpurun_out/bench_c4_r5a.err
python -c "
import json; d=json.loads(open('gpurun_out/bench_c4_r5a.json').read().splitlines()[-1]); print(d['value'], 'e2e', d['e2e']['value'], json.dumps(d['e2e']['last_call_ms_rank0'])); print('   arrays', d['e2e_arrays']['value'], d['e2e_arrays']['last_call_ms_rank0']['ms_call'])"
