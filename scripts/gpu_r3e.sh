#!/bin/bash
# Newick parser round: tests, standalone parser timing, bench with / without the CTA slot reserved for the parser
mkdir -p gpurun_out
python -m pytest tests/test_gpu_newick.py -x -q 2>&1 | tail -30 > gpurun_out/gputests_r3e.log; cat gpurun_out/gputests_r3e.log
python scripts/gpu_newick_probe.py c4 16384 2>&1 | tail -8
python scripts/gpu_newick_probe.py c3 4096 2>&1 | tail -8
for r in 1 0; do
  python bench.py --no-extras --no-cpu-baseline --opt parse_reserve=$r > gpurun_out/bench_c4_r3e_$r.json 2> gpurun_out/bench_c4_r3e.err; tail -3 gpurun_out/bench_c4_r3e.err
  python -c "
import json; d=json.load(open('gpurun_out/bench_c4_r3e_$r.json')); print(d['value'], d['e2e']['value'], json.dumps(d['e2e']['last_call_ms_rank0'])); print(d['e2e_arrays']['value'])"
done
