#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests/test_gpu_newick.py tests/test_scaling_regime.py -x -q -m gpu 2>&1 | tail -30 > gpurun_out/gputests_r3f.log; cat gpurun_out/gputests_r3f.log
python scripts/gpu_newick_probe.py c4 16384 2>&1 | grep '"kernel": 2' | tail -2
python scripts/gpu_newick_probe.py c3 4096 2>&1 | grep '"kernel": 2' | tail -2
for r in 1 0; do
  python bench.py --no-extras --no-cpu-baseline --opt parse_reserve=$r > gpurun_out/bench_c4_r3f_$r.json 2> gpurun_out/bench_c4_r3f.err; tail -3 gpurun_out/bench_c4_r3f.err
  python -c "
import json; d=json.load(open('gpurun_out/bench_c4_r3f_$r.json')); print(d['value'], d['e2e']['value'], json.dumps(d['e2e']['last_call_ms_rank0'])); print(d['e2e_arrays']['value'])"
done
python bench.py --config c3 --no-cpu-baseline > gpurun_out/bench_c3_r3f.json 2>> gpurun_out/bench_c4_r3f.err
python -c "
import json; d=json.load(open('gpurun_out/bench_c3_r3f.json')); print(d['value'], d['e2e']['value'], json.dumps(d['e2e']['last_call_ms_rank0'])); print(d['e2e_arrays']['value'])"
