#!/bin/bash
# e2e sweep of the pipelined call's chunk size (waves per launch); writes gpurun_out/sweep_<tag>.jsonl
tag=${1:-sw}
mkdir -p gpurun_out
: > gpurun_out/sweep_$tag.jsonl
for cfg in c4 c3; do
  for w in 1 2 3 4 6; do
    python bench.py --config $cfg --no-cpu-baseline --opt pipe_waves=$w >> gpurun_out/sweep_$tag.jsonl 2>> gpurun_out/sweep_$tag.err
  done
done
python - "$tag" <<'PY'
import json,sys
for l in open("gpurun_out/sweep_%s.jsonl" % sys.argv[1]):
    d=json.loads(l); print(d["config"]["workload"][:2], d["config"].get("options"), round(d["value"]), round(d["e2e"]["value"]), d["e2e"]["last_call_ms_rank0"])
PY
