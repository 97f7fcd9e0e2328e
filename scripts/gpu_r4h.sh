#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests/test_gpu_phylohmm.py tests/test_gpu_group.py -x -q -m gpu 2>&1 | tail -12
LH_TRACE=1 python bench.py --config c3 > gpurun_out/bench_c3_r4h.json 2> gpurun_out/bench_c3_r4h.err; echo rc=$?
grep "RunPipeline" gpurun_out/bench_c3_r4h.err | tail -3
python - <<PY
import json
d = json.loads(open("gpurun_out/bench_c3_r4h.json").read().splitlines()[-1])
print("c3 value", d["value"], "e2e", d["e2e"]["value"], "pipeline", json.dumps(d["e2e_pipeline"])[:700])
PY
