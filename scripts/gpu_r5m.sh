#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests/test_gpu_newick.py tests/test_gpu_phylohmm.py tests/test_gpu_group.py -x -q -m gpu 2>&1 | tail -3
for flag in "" "--pageable-text"; do
  python bench.py --no-extras --no-cpu-baseline $flag > gpurun_out/bench_c4_r5m.json 2> gpurun_out/bench_c4_r5m.err; tail -1 gpurun_out/bench_c4_r5m.err | cut -c1-200
  python -c "
import json; d=json.loads(open('gpurun_out/bench_c4_r5m.json').read().splitlines()[-1]); print('[$flag]', d['value'], 'e2e', d['e2e']['value'], d['e2e']['text_memory'], json.dumps(d['e2e']['last_call_ms_rank0']))"
  LH_TRACE=1 taskset -c 0-3 python bench.py --batch 2048 --opt host_threads=4 --no-extras --no-cpu-baseline --steps 3 --warmup 3 $flag > gpurun_out/trace_r5m.json 2> gpurun_out/trace_r5m.err
  python -c "
import json; d=json.loads(open('gpurun_out/trace_r5m.json').read().splitlines()[-1]); print('   1-of-8 rank [$flag]', 'e2e', d['e2e']['value'], json.dumps(d['e2e']['last_call_ms_rank0']))"
  grep "dev compute chunk" gpurun_out/trace_r5m.err | tail -10 | tr '\n' ';'; echo
done
