#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests/test_gpu_asr.py -x -q -m gpu 2>&1 | tail -2
python bench_asr.py --config c3 > gpurun_out/bench_asr_c3_r5l.json 2> gpurun_out/bench_asr_c3.err; tail -2 gpurun_out/bench_asr_c3.err; cut -c1-330 gpurun_out/bench_asr_c3_r5l.json
python bench_asr.py --config c4 > gpurun_out/bench_asr_c4_r5l.json 2> gpurun_out/bench_asr_c4.err; tail -2 gpurun_out/bench_asr_c4.err; cut -c1-330 gpurun_out/bench_asr_c4_r5l.json
