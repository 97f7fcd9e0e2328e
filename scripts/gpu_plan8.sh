#!/bin/bash
# device planner at 8 ranks: C4 with k_plan forced on / off / auto
tag=${1:-rX}
mkdir -p gpurun_out
for dp in 1 0 -1; do
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port $((29600 + dp + 1)) \
    bench.py --gpus 8 --no-cpu-baseline --opt device_plan=$dp > gpurun_out/bench_c4_8gpu_dp${dp}_$tag.json 2> gpurun_out/bench_c4_8gpu_dp${dp}_$tag.err
  tail -c 1300 gpurun_out/bench_c4_8gpu_dp${dp}_$tag.json; echo
done
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29650 \
    bench.py --gpus 8 --config c5 --families 1024 --no-cpu-baseline > gpurun_out/bench_c5_8gpu_$tag.json 2> gpurun_out/bench_c5_8gpu_$tag.err
tail -c 1500 gpurun_out/bench_c5_8gpu_$tag.json; tail -3 gpurun_out/bench_c5_8gpu_$tag.err
nproc
