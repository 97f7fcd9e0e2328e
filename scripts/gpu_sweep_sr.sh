#!/bin/bash
# k_prune_sr arena sweep: occupancy vs overflow to the global part
tag=${1:-sr}
: > gpurun_out/sweep_$tag.jsonl
for cfg in c4 c3; do
  for a in 64 256 512; do
    python bench.py --config $cfg --no-cpu-baseline --no-e2e --steps 2 --opt site_repeats=1 --opt sr_arena=$a >> gpurun_out/sweep_$tag.jsonl 2>> gpurun_out/sweep_$tag.err
  done
done
python - "$tag" <<'PY'
import json,sys
for l in open("gpurun_out/sweep_%s.jsonl" % sys.argv[1]):
    d=json.loads(l); print(d["config"]["workload"][:2], d["config"].get("options"), d["config"].get("sr_ctas_per_sm"), round(d["value"]), {k: round(v,2) for k,v in d["stage_ms_per_step"].items() if k!="rank"})
PY
