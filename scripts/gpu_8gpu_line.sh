#!/bin/bash
# the default bench line under torchrun at N ranks (what the driver runs for SCALE)
mkdir -p gpurun_out
n=${1:-8}; tag=${2:-rX}
python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus $n > gpurun_out/bench_c4_${n}gpu_$tag.json 2> gpurun_out/bench_c4_${n}gpu_$tag.err
echo rc=$?; tail -2 gpurun_out/bench_c4_${n}gpu_$tag.err
python - <<PY
import json
d = json.loads(open("gpurun_out/bench_c4_${n}gpu_$tag.json").read().splitlines()[-1])
print("value", d["value"], "ms/step", d["ms_per_step"], d["stage_ms_per_step"])
print("e2e", d["e2e"]["value"], d["e2e"]["last_call_ms_rank0"])
print("arrays", d["e2e_arrays"]["value"], d["e2e_arrays"]["last_call_ms_rank0"])
print("parity_vs_n1", d["parity_vs_n1"]["equal"], "c3", d["c3"]["value"], d["c3"]["e2e"]["value"], d["c3"]["parity_vs_n1"]["equal"], "c5", d["c5"]["value"], d["c5"]["parity"]["max_rel_err_vs_cpu"])
PY
