#!/bin/bash
LH_TRACE=1 python scripts/gpu_pipe_probe.py 4096 2> gpurun_out/pipe_trace.err | grep "batch 256"
python - <<'PY'
lines=[l.rstrip() for l in open('gpurun_out/pipe_trace.err') if l.startswith('LHTRACE')]
idx=[i for i,l in enumerate(lines) if l.split()[1]=='check']
calls=[lines[a:b] for a,b in zip(idx, idx[1:]+[len(lines)])]
# a steady-state call in the middle of the 256-batch runs
c=calls[len(calls)//3]
print("\n".join(c[:60]))
PY
