#!/bin/bash
# eight-GPU round: BASELINE config 5 in full (10 000 families x 1000 trees) and the default bench line at N=8
mkdir -p gpurun_out
tag=${1:-rX}
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1"
$TR --master-port 29521 bench.py --gpus 8 --config c5 --families 10000 --steps 2 --warmup 3 > gpurun_out/bench_c5_10k_8gpu_$tag.json 2> gpurun_out/bench_c5_10k_8gpu_$tag.err
echo rc=$?; tail -2 gpurun_out/bench_c5_10k_8gpu_$tag.err; tail -c 1500 gpurun_out/bench_c5_10k_8gpu_$tag.json
$TR --master-port 29522 bench.py --gpus 8 > gpurun_out/bench_c4_8gpu_$tag.json 2> gpurun_out/bench_c4_8gpu_$tag.err
echo rc=$?; tail -2 gpurun_out/bench_c4_8gpu_$tag.err
python - <<PY
import json
d = json.loads(open("gpurun_out/bench_c4_8gpu_$tag.json").read().splitlines()[-1])
print("value", d["value"], "e2e", d["e2e"]["value"], "arrays", d["e2e_arrays"]["value"], "parity_vs_n1", d["parity_vs_n1"])
print("c3", d["c3"]["value"], d["c3"]["e2e"]["value"], d["c3"]["parity_vs_n1"])
print("c5", d["c5"]["value"], d["c5"]["parity"])
PY
python -m pytest tests/test_gpu_group.py -x -q -m gpu 2>&1 | tail -3
