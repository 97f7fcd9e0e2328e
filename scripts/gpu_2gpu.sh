#!/bin/bash
# two-GPU round: the group tests and the bench under torchrun at N=2
mkdir -p gpurun_out
tag=${1:-rX}
python -m pytest tests/test_gpu_group.py -x -q -m gpu 2>&1 | tail -12 > gpurun_out/gputests_group_$tag.log; cat gpurun_out/gputests_group_$tag.log
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 3 --warmup 3 > gpurun_out/bench_c4_2gpu_$tag.json 2> gpurun_out/bench_c4_2gpu_$tag.err
echo rc=$?; tail -5 gpurun_out/bench_c4_2gpu_$tag.err
python - <<PY
import json
d = json.load(open("gpurun_out/bench_c4_2gpu_$tag.json"))
print("value", d["value"], "e2e", d["e2e"]["value"], "arrays", d["e2e_arrays"]["value"], "parity_vs_n1", d["parity_vs_n1"], "parity", d["parity"])
print("c3", d["c3"]["value"], d["c3"]["e2e"]["value"], d["c3"]["parity_vs_n1"], d["c3"]["parity"])
print("c5", d["c5"]["value"], d["c5"]["parity"])
PY
