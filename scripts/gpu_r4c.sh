#!/bin/bash
mkdir -p gpurun_out
tag=${1:-r4c}
python -m pytest tests -m gpu -x -q 2>&1 | tail -12 > gpurun_out/gputests_$tag.log; cat gpurun_out/gputests_$tag.log
python bench.py > gpurun_out/bench_c4_$tag.json 2> gpurun_out/bench_c4_$tag.err; echo rc=$?; tail -3 gpurun_out/bench_c4_$tag.err
python bench.py --impl reference > gpurun_out/bench_ref_c4_$tag.json 2>> gpurun_out/bench_c4_$tag.err
python - <<PY
import json
d = json.loads(open("gpurun_out/bench_c4_$tag.json").read().splitlines()[-1])
r = json.loads(open("gpurun_out/bench_ref_c4_$tag.json").read().splitlines()[-1])
print("c4 value", d["value"], "e2e", d["e2e"]["value"], "arrays", d["e2e_arrays"]["value"], "roofline", d["roofline"]["frac"], "stage", d["stage_ms_per_step"])
print("ref arm", r["value"], "cpu_baseline", {k: (v["value"] if isinstance(v, dict) else v) for k, v in d["cpu_baseline"].items() if k in ("value", "single_thread", "as_shipped")})
print("c3 value", d["c3"]["value"], "e2e", d["c3"]["e2e"]["value"], "arrays", d["c3"]["e2e_arrays"]["value"])
print("c3 pipeline", json.dumps(d["c3"]["e2e_pipeline"]))
print("c5", d["c5"]["value"], d["c5"]["parity"]["max_rel_err_vs_cpu"])
PY
