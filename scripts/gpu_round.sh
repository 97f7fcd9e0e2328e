#!/bin/bash
# One GPU-box round: parity tests, the bench lines, the ncu launch list and one full capture of k_prune.
# usage: scripts/gpu_round.sh <tag>      (run through gpurun; writes gpurun_out/*_<tag>.*)
tag=${1:-rX}
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q 2>&1 | tail -5 > gpurun_out/gputests_$tag.log; cat gpurun_out/gputests_$tag.log
python bench.py > gpurun_out/bench_c4_$tag.json 2> gpurun_out/bench_c4_$tag.err; tail -c 3000 gpurun_out/bench_c4_$tag.json; tail -3 gpurun_out/bench_c4_$tag.err
python bench.py --impl reference > gpurun_out/bench_ref_c4_$tag.json 2>> gpurun_out/bench_c4_$tag.err; cat gpurun_out/bench_ref_c4_$tag.json
python bench.py --config c3 > gpurun_out/bench_c3_$tag.json 2> gpurun_out/bench_c3_$tag.err; tail -c 3000 gpurun_out/bench_c3_$tag.json
if [ "$2" != "noncu" ]; then
python bench.py --no-cpu-baseline > gpurun_out/plain_$tag.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file gpurun_out/launches_$tag.csv python bench.py --no-cpu-baseline > gpurun_out/ncu_launch_$tag.log 2>&1
python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-e2e --batch 4096 > gpurun_out/plain2_$tag.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:k_prune -s 3 -c 1 -o gpurun_out/prof_prune_$tag -f python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-e2e --batch 4096 > gpurun_out/ncu_full_$tag.log 2>&1
tail -3 gpurun_out/ncu_full_$tag.log
fi
