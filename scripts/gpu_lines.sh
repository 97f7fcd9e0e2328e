#!/bin/bash
# bench lines of every config on one GPU (no ncu)
tag=${1:-rX}
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q 2>&1 | tail -3
for c in c4 c3 c2 c5; do
  python bench.py --config $c > gpurun_out/bench_${c}_$tag.json 2> gpurun_out/bench_${c}_$tag.err
  python - <<PY
import json
d = json.load(open("gpurun_out/bench_${c}_$tag.json"))
print("$c", "value", round(d["value"]), "e2e", round(d["e2e"]["value"]), "newick", d.get("e2e_newick") and round(d["e2e_newick"]["value"]), "parity", d.get("parity"))
PY
done
