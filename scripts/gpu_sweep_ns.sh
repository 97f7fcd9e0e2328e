#!/bin/bash
tag=${1:-ns}
: > gpurun_out/sweep_$tag.jsonl
for cfg in c4 c3; do
  for ns in 4 2 1; do
    python bench.py --config $cfg --no-cpu-baseline --no-e2e --steps 3 --opt sites_per_thread=$ns >> gpurun_out/sweep_$tag.jsonl 2>> gpurun_out/sweep_$tag.err
  done
done
python - "$tag" <<'PY'
import json,sys
for l in open("gpurun_out/sweep_%s.jsonl" % sys.argv[1]):
    d=json.loads(l); print(d["config"]["workload"][:2], d["config"].get("options"), round(d["value"]), {k: round(v,2) for k,v in d["stage_ms_per_step"].items() if k!="rank"})
PY
