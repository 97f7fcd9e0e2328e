#!/bin/bash
# what the driver runs at round end: smoke, the gpu suite, the default bench line and the reference arm
mkdir -p gpurun_out
tag=${1:-final}
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -3
python -m pytest tests -m gpu -x -q 2>&1 | tail -4
python bench.py > gpurun_out/bench_default_$tag.json 2> gpurun_out/bench_default_$tag.err; echo "bench rc=$?"; tail -2 gpurun_out/bench_default_$tag.err | cut -c1-300
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_reference_$tag.json 2> gpurun_out/bench_reference_$tag.err; echo "ref rc=$?"; cut -c1-400 gpurun_out/bench_reference_$tag.json
python - <<PY
import json
d = json.loads(open("gpurun_out/bench_default_$tag.json").read().splitlines()[-1])
print("value", d["value"], "e2e", d["e2e"]["value"], d["e2e"]["text_memory"], "arrays", d["e2e_arrays"]["value"], "launches", d["gpu_launches"])
print("roofline", d["roofline"]["frac"], d["roofline"]["bound"], "cpu", {k: (v if not isinstance(v, dict) else v.get("value")) for k, v in d["cpu_baseline"].items() if k in ("value", "single_thread", "as_shipped", "kind", "cores")})
print("clocks", d["clocks"]); print("c3", d["c3"]["value"], d["c3"]["e2e"]["value"], d["c3"].get("e2e_pipeline"), "c5", d["c5"]["value"])
PY
