#!/bin/bash
# Strong-scaling lines of bench.py at N = 2, 4, 8 on one box (run through gpurun --gpus 8).
tag=${1:-rX}
mkdir -p gpurun_out
for n in 8 4 2; do
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29500 + n)) \
    bench.py --gpus $n > gpurun_out/bench_c4_${n}gpu_$tag.json 2> gpurun_out/bench_c4_${n}gpu_$tag.err
  python - <<PY
import json
try:
    d = json.loads(open("gpurun_out/bench_c4_${n}gpu_$tag.json").read().strip().splitlines()[-1])
    print($n, "gpus value", d["value"], "e2e", d["e2e"]["value"], "newick", d["e2e_newick"]["value"], d["clocks"])
except Exception as e:
    print($n, "failed", e)
PY
  tail -2 gpurun_out/bench_c4_${n}gpu_$tag.err
done
