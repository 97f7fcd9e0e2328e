#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests/test_gpu_newick.py tests/test_zz_no_alloc.py -x -q -m gpu 2>&1 | tail -3
python scripts/gpu_newick_probe.py c4 16384 2>&1 | grep '"kernel": 2' | tail -1
python scripts/gpu_newick_probe.py c3 4096 2>&1 | grep '"kernel": 2' | tail -1
python bench.py --no-extras --no-cpu-baseline > gpurun_out/bench_c4_r4g.json 2> gpurun_out/bench_c4_r4g.err; tail -2 gpurun_out/bench_c4_r4g.err
python -c "
import json; d=json.loads(open('gpurun_out/bench_c4_r4g.json').read().splitlines()[-1]); print(d['value'], d['e2e']['value'], json.dumps(d['e2e']['last_call_ms_rank0'])); print(d['e2e_arrays']['value'])"
