#!/bin/bash
# the timeline of one rank's text-in call at the 8-GPU shard size (2048 trees, 4 host threads), traced
mkdir -p gpurun_out
LH_TRACE=1 taskset -c 0-3 python bench.py --batch 2048 --opt host_threads=4 --no-extras --no-cpu-baseline --steps 3 --warmup 3 > gpurun_out/trace_r5f.json 2> gpurun_out/trace_r5f.err
python - <<'PY'
import json,re
d=json.loads(open('gpurun_out/trace_r5f.json').read().splitlines()[-1]); print(d['value'], 'e2e', d['e2e']['value'], json.dumps(d['e2e']['last_call_ms_rank0'])); print('   arrays', d['e2e_arrays']['value'], d['e2e_arrays']['last_call_ms_rank0']['ms_call'])
lines=[l for l in open('gpurun_out/trace_r5f.err') if l.startswith('LHTRACE')]
# calls are separated by "check" marks; print the last two newick calls
idx=[i for i,l in enumerate(lines) if l.split()[1]=='check']
print(len(lines), len(idx))
calls=[lines[a:b] for a,b in zip(idx, idx[1:]+[len(lines)])]
nw=[c for c in calls if any('text' in l or 'newick' in l or 'parse' in l for l in c)]
for c in (nw[-1:] if nw else calls[-1:]):
    print(''.join(c[:120]))
PY
