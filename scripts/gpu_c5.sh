#!/bin/bash
# repertoire (config c5) round: tests, slot sweep
tag=${1:-rX}
mkdir -p gpurun_out
python -m pytest tests/test_gpu_repertoire.py -x -q 2>&1 | tail -15 > gpurun_out/gputests_c5_$tag.log; cat gpurun_out/gputests_c5_$tag.log
for slots in 1 2 3 4; do
python bench.py --config c5 --slots $slots --no-cpu-baseline > gpurun_out/bench_c5_s${slots}_$tag.json 2> gpurun_out/bench_c5_$tag.err; tail -c 1500 gpurun_out/bench_c5_s${slots}_$tag.json; tail -3 gpurun_out/bench_c5_$tag.err
done
python bench.py --config c5 > gpurun_out/bench_c5_$tag.json 2>> gpurun_out/bench_c5_$tag.err; tail -c 2500 gpurun_out/bench_c5_$tag.json
