#!/bin/bash
mkdir -p gpurun_out
python bench_asr.py --config c3 > gpurun_out/bench_asr_c3_r5k.json 2> gpurun_out/bench_asr_c3.err; tail -2 gpurun_out/bench_asr_c3.err; cat gpurun_out/bench_asr_c3_r5k.json | cut -c1-900
python bench_asr.py --config c4 > gpurun_out/bench_asr_c4_r5k.json 2> gpurun_out/bench_asr_c4.err; tail -2 gpurun_out/bench_asr_c4.err; cat gpurun_out/bench_asr_c4_r5k.json | cut -c1-900
ncu --set full --clock-control none --import-source on -k regex:k_asr$ -s 2 -c 1 -o gpurun_out/prof_k_asr_r5k -f python bench_asr.py --config c3 --no-cpu --steps 1 --warmup 2 > gpurun_out/ncu_asr.log 2>&1; ls -la gpurun_out/prof_k_asr_r5k.ncu-rep
